// First-layer convolution (1 -> 64 channels, 5x5x5, stride 2, no padding) with ZERO im2col work.
//
// Space-to-depth view: a stride-2 5^3 convolution over a 1-channel volume is a stride-1 3^3 convolution over the
// 8-channel "parity" volume  s2d[z'][y'][x'][c = (pz,py,px)] = x[2z'+pz][2y'+py][2x'+px]  (taps with k = 5 get weight 0).
// One s2d position is 8 bf16 = 16 bytes = exactly one UMMA core-matrix row of the no-swizzle K-major layout, and
// consecutive output columns are consecutive s2d positions.  So a shared-memory line  line[x'][8ch]  (pitch 64
// positions = 1 KiB) can be handed to tcgen05.mma directly: the descriptor's row stride is the 16-byte position
// stride, and the K chunk stride (LBO) is ALSO 16 bytes -- chunk 1 of row r is row r+1's data, which is precisely the
// shifted tap (jx+1).  Lines of consecutive y' are consecutive in memory, so rows 64..127 of the M=128 tile are the
// next output line.  Per 2x59-output tile: 18 MMAs read 12 staged lines; no thread ever gathers an im2col row.
//   fprop : D[pos][co]      = sum_{jd,jh,jw} line(jd,jh)[pos+jw][c] * Ws2d[co][(jd,jh,jw),c]
//   wgrad : D[co][jw*8+c]   = sum_pos dY[pos][co] * line(jd,jh)[pos+jw][c]      (9 accumulators, one per (jd,jh))
// The staging warps only convert the 8-bit volume (u8 * scale -> bf16) into parity-interleaved lines.
// Reference op: nn.Conv3d(1,64,5,stride=2) + its weight gradient, fedml_api/model/cv/salient_models.py:147 (cuDNN fp32).
#include "tc_common.cuh"
#include <stdlib.h>

namespace nidt {
namespace s2d {

struct Geom {
    int N, Di, Hi, Wi;     // u8 input volume
    int Do, Ho, Wo;        // output extents ((Di-5)/2+1 ...), Wo <= 62
    int h_tiles;           // ceil(Ho / 2): a tile is 2 output lines
    int total_tiles;       // N * Do * h_tiles
    int raw_pitch;         // bytes per plane of the raw staging buffer (8 input rows + misalignment, multiple of 16)
    long long in_bytes;    // size of the input tensor (bulk copies never read past its 16-byte-rounded end)
    float scale;
    int dbg;               // profiling aid (NIDT_S2D_DEBUG): bit0 skip line building, bit1 skip MMAs, bit2 skip dY loads, bit3 skip raw copies
};

constexpr int kThreads = 13 * 32;        // wgrad: warps 0-3 / 4-7 staging groups, warp 8 MMA, warps 9-12 producers (dY TMA, raw bulk copies) + final epilogue
constexpr int kFpThreads = 18 * 32;      // fprop: ... warps 9-12 / 13-16 = two epilogue groups (tile i -> group i & 1), warp 17 = raw producer
constexpr int kLineBytes = 64 * 16;      // one s2d line: 64 positions x 8 bf16
constexpr int kLines = 12;               // 3 jd planes x 4 y' lines
constexpr int kABytes = 16 * kLineBytes + 2048 + 256;   // 12 lines are written; wgrad addresses 16 line-chunks x 128 positions (+ taps): the rest stays zero
constexpr int kPlanes = 6, kRows = 8;    // raw input planes / rows per tile
constexpr int kT16 = 16384;
constexpr int kRawRing = 4;               // raw u8 tiles in flight (bulk copies issued by one producer thread)
constexpr int kARing = 4;                 // line buffers: each staging group runs one of its own tiles ahead of the MMAs
constexpr int kWgPitch = 65;              // wgrad: positions per line.  65 (not 64) skews consecutive lines by 16 B so that the 16 line-chunks of one
                                          // MN-major A row fall into different shared-memory banks (a 1 KiB pitch would be a 16-way conflict)
constexpr int kWgLineBytes = kWgPitch * 16;
constexpr int kWgABytes = 16 * kWgLineBytes + 2048 + 256;
constexpr int kDyRing = 4;                // dY tiles in flight ahead of the wgrad MMAs

__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// no-swizzle descriptor (layout type 0): LBO = stride between the two K chunks (K-major) / between K groups (MN-major),
// SBO = stride between 8-row groups (K-major) / between MN chunks (MN-major)
__device__ __forceinline__ uint64_t desc_noswz(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}

// Tile walker without per-tile integer divisions: tiles advance by a fixed stride, so (ht, d, n) advance by fixed deltas with carries.
struct TileIter {
    int tile, ht, d, n;
    int s_ht, s_d, s_n, stride;
    __device__ __forceinline__ TileIter(const Geom& g, int first, int stride_) : tile(first), stride(stride_) {
        ht = first % g.h_tiles;
        const int t2 = first / g.h_tiles;
        d = t2 % g.Do;
        n = t2 / g.Do;
        s_ht = stride_ % g.h_tiles;
        const int q = stride_ / g.h_tiles;
        s_d = q % g.Do;
        s_n = q / g.Do;
    }
    __device__ __forceinline__ bool valid(const Geom& g) const { return tile < g.total_tiles; }
    __device__ __forceinline__ int h0() const { return ht * 2; }
    __device__ __forceinline__ int lines(const Geom& g) const { return min(2, g.Ho - ht * 2); }
    __device__ __forceinline__ void next(const Geom& g) {
        tile += stride;
        ht += s_ht;
        int c = 0;
        if (ht >= g.h_tiles) { ht -= g.h_tiles; c = 1; }
        d += s_d + c;
        c = 0;
        if (d >= g.Do) { d -= g.Do; c = 1; }
        n += s_n + c;
    }
};

__device__ __forceinline__ void decode_tile(const Geom& g, int tile, int& n, int& d, int& h0, int& lines) {
    const int ht = tile % g.h_tiles;
    const int t2 = tile / g.h_tiles;
    d = t2 % g.Do;
    n = t2 / g.Do;
    h0 = ht * 2;
    lines = min(2, g.Ho - h0);
}

__device__ __forceinline__ const uint8_t* plane_src(const uint8_t* __restrict__ x, const Geom& g, int n, int z, int y0) {
    return x + (((size_t)n * g.Di + z) * g.Hi + y0) * g.Wi;
}

// producer thread: the 8 input rows of each of the 6 planes are one contiguous byte run per plane -> one bulk copy per plane
// (aligned superset of the run) into raw slot `tile_seq % kRawRing`
__device__ __forceinline__ void raw_producer_loop(const uint8_t* __restrict__ x, const Geom& g, uint8_t* raws, uint64_t* raw_full, uint64_t* raw_empty) {
    const uintptr_t x_end = (reinterpret_cast<uintptr_t>(x) + (uintptr_t)g.in_bytes + 15) & ~(uintptr_t)15;
    TileIter t(g, blockIdx.x, gridDim.x);
    for (int i = 0; t.valid(g); ++i, t.next(g)) {
        const int slot = i % kRawRing;
        const int n = t.n;
        tc::mbar_wait(&raw_empty[slot], ((i / kRawRing) & 1) ^ 1);
        const int z0 = 2 * t.d, y0 = 2 * t.h0();
        const int planes = min(kPlanes, g.Di - z0);
        uint32_t total = 0, sz[kPlanes];
        const uint8_t* src[kPlanes];
#pragma unroll
        for (int p = 0; p < kPlanes; ++p) {
            sz[p] = 0;
            if (p < planes) {
                const uint8_t* s0 = plane_src(x, g, n, z0 + p, y0);
                src[p] = s0 - (reinterpret_cast<uintptr_t>(s0) & 15);
                const uintptr_t room = x_end - reinterpret_cast<uintptr_t>(src[p]);
                sz[p] = (uint32_t)min((uintptr_t)g.raw_pitch, room);
                total += sz[p];
            }
        }
        if (g.dbg & 8) { tc::mbar_arrive(&raw_full[slot]); continue; }
        tc::mbar_expect_tx(&raw_full[slot], total);
#pragma unroll
        for (int p = 0; p < kPlanes; ++p)
            if (sz[p]) tc::bulk_load_1d(raws + (slot * kPlanes + p) * g.raw_pitch, src[p], sz[p], &raw_full[slot]);
    }
}

// raw u8 rows -> 12 parity-interleaved bf16 lines.  One item = TWO adjacent s2d positions (32 bytes): for each of the
// 4 (pd,ph) input rows one unaligned 4-byte read (two aligned words + funnel shift) yields x[4i .. 4i+3]; bytes become
// floats without the slow I2F path: float(0x4B000000 | b) - 2^23 = b exactly (one PRMT + one FADD), then the scale multiply.
// A thread always handles the same three items, so everything that does not depend on the tile is computed once (LineCtx).
struct LineCtx {
    uint32_t base[3][4];     // plane*raw_pitch + row*Wi + 4*i4   per (item, pp)
    uint32_t dst[3];         // byte offset of the item inside a line buffer
    uint32_t bmask[3];       // valid bytes of the 4-byte group (row end)
    uint32_t psz;            // plane-to-plane change of the 16-byte misalignment
    __device__ __forceinline__ LineCtx(const Geom& g, int gt, int line_bytes) {
        psz = (uint32_t)((size_t)g.Hi * g.Wi) & 15u;
#pragma unroll
        for (int it = 0; it < 3; ++it) {
            const int item = gt + it * 128;
            const int i4 = item & 31, L = item >> 5, jd = L >> 2, hl = L & 3;
            const int nvalid = min(4, max(0, g.Wi - 4 * i4));
            bmask[it] = nvalid >= 4 ? 0xffffffffu : ((1u << (8 * nvalid)) - 1u);
            dst[it] = L * line_bytes + i4 * 32;
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) base[it][pp] = (2 * jd + (pp >> 1)) * g.raw_pitch + (2 * hl + (pp & 1)) * g.Wi + 4 * i4;
        }
    }
};

__device__ __forceinline__ void build_lines(const LineCtx& c, uint32_t mis0, int rows_ok, int planes_ok, const uint8_t* __restrict__ raw,
                                            uint8_t* __restrict__ a_buf, float scale, int gt) {
#pragma unroll
    for (int it = 0; it < 3; ++it) {                       // fully unrolled: three independent load->convert->store chains in flight
        const int L = (gt + it * 128) >> 5, jd = L >> 2, hl = L & 3;
        uint32_t lo[4], hi[4];
#pragma unroll
        for (int pp = 0; pp < 4; ++pp) {                   // pp = pd*2 + ph  -> channels (pd,ph,0), (pd,ph,1)
            const int plane = 2 * jd + (pp >> 1), row = 2 * hl + (pp & 1);
            uint32_t four = 0;
            if (plane < planes_ok && row < rows_ok) {
                const uint32_t a = c.base[it][pp] + ((mis0 + plane * c.psz) & 15u);
                const uint32_t* rw = reinterpret_cast<const uint32_t*>(raw + (a & ~3u));
                four = __funnelshift_r(rw[0], rw[1], (a & 3u) * 8) & c.bmask[it];
            }
            const float f0 = (__uint_as_float(__byte_perm(four, 0x4B000000u, 0x7650)) - 8388608.f) * scale;
            const float f1 = (__uint_as_float(__byte_perm(four, 0x4B000000u, 0x7651)) - 8388608.f) * scale;
            const float f2 = (__uint_as_float(__byte_perm(four, 0x4B000000u, 0x7652)) - 8388608.f) * scale;
            const float f3 = (__uint_as_float(__byte_perm(four, 0x4B000000u, 0x7653)) - 8388608.f) * scale;
            __nv_bfloat162 p0 = __floats2bfloat162_rn(f0, f1), p1 = __floats2bfloat162_rn(f2, f3);
            lo[pp] = *reinterpret_cast<uint32_t*>(&p0);
            hi[pp] = *reinterpret_cast<uint32_t*>(&p1);
        }
        uint4* dst = reinterpret_cast<uint4*>(a_buf + c.dst[it]);
        dst[0] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        dst[1] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    }
}

// staging warps (0-7): two groups alternate tiles.  Raw bytes arrive by bulk copy kRawRing tiles ahead; a group converts
// them into the line buffer `i % kARing` as soon as the MMAs that read that buffer kARing tiles ago have finished.
__device__ __forceinline__ void staging_loop(const uint8_t* __restrict__ x, const Geom& g, uint8_t* a_bufs, uint8_t* raws, uint64_t* a_full, uint64_t* a_empty,
                                             uint64_t* raw_full, uint64_t* raw_empty, int line_bytes, int a_bytes) {
    const int grp = threadIdx.x >> 7, gt = threadIdx.x & 127;
    const LineCtx ctx(g, gt, line_bytes);
    TileIter t(g, blockIdx.x + grp * gridDim.x, 2 * gridDim.x);
    for (int i = grp; t.valid(g); i += 2, t.next(g)) {
        const int buf = i % kARing, slot = i % kRawRing;
        const int z0 = 2 * t.d, y0 = 2 * t.h0();
        const uint32_t mis0 = (uint32_t)(reinterpret_cast<uintptr_t>(plane_src(x, g, t.n, z0, y0)) & 15);
        tc::mbar_wait(&raw_full[slot], (i / kRawRing) & 1);
        tc::mbar_wait(&a_empty[buf], ((i / kARing) & 1) ^ 1);
        if (!(g.dbg & 1)) build_lines(ctx, mis0, min(kRows, g.Hi - y0), min(kPlanes, g.Di - z0), raws + slot * kPlanes * g.raw_pitch, a_bufs + buf * a_bytes, g.scale, gt);
        if (!(g.dbg & 16)) tc::fence_proxy_async_smem();
        __syncwarp();
        if ((threadIdx.x & 31) == 0) {                   // one arrival per warp (128 same-address shared atomics per tile were a bottleneck)
            tc::mbar_arrive(&a_full[buf]);
            tc::mbar_arrive(&raw_empty[slot]);
        }
    }
}

// ======================================================================================= fprop
// Optional epilogue for consumers that only need the layer's output max-pooled 3x3x3 (the engine's forward when the backward does not
// read y): instead of storing the tile, the epilogue reduces every 3 consecutive output columns of a line to the one BatchNorm + ReLU
// + MaxPool would pick -- the maximum of y where gamma >= 0, the minimum where gamma < 0 (the affine map is monotone per channel;
// first one on ties) -- and stores that value (raw y, bf16) plus its column 0..2.  A third of the bytes written, and the pooling
// kernel that follows (nidt_bn_relu_pool3w_fwd) reads 9 candidates per window instead of 27.  BatchNorm statistics are unaffected
// (they are summed from the staged tile either way).
struct PoolW {
    __nv_bfloat16* yw;          // [N][Do][Ho][Wp][64]; nullptr = store the full tile through tm_y
    unsigned short* aw;         // [N][Do][Ho][Wp][8]: 2 bits per channel (column of the winner) for the 8 channels of a 16-byte chunk
    const float* gamma;         // BatchNorm weight: its sign decides max / min
    int Wp;                     // (Wo - 3) / 3 + 1
};

__global__ void __launch_bounds__(kFpThreads, 1) fprop_kernel(const uint8_t* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                                              const __grid_constant__ CUtensorMap tm_y, float* __restrict__ stats_partials, const Geom g,
                                                              const PoolW pw) {
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint8_t* staging = smem;                              // [2][128 rows][128 B] swizzled (TMA store source), one per epilogue group
    uint8_t* b_tile = staging + 2 * kT16;                 // [36 K-chunks][64 co][16 B]  (no swizzle)
    uint8_t* a_bufs = b_tile + 36 * 1024;                 // [kARing][kABytes]
    uint64_t* bars = reinterpret_cast<uint64_t*>(a_bufs + kARing * kABytes);
    uint64_t *a_full = bars, *a_empty = bars + kARing, *t_full = bars + 2 * kARing, *t_empty = t_full + 2;
    uint64_t *raw_full = t_empty + 2, *raw_empty = raw_full + kRawRing;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(raw_empty + kRawRing);
    float* sbias = reinterpret_cast<float*>(tmem_ptr + 4);    // [64]
    float* red = sbias + 64;                                   // [2 groups][16 parts][2][64]
    uint8_t* raws = reinterpret_cast<uint8_t*>(red + 2 * 16 * 2 * 64);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler: no per-MMA lane-election loops

    // s2d weights: chunk kc = ((jd*3+jh)*4 + jw), channel c = pd*4+ph*2+pw  ->  W[co][2jd+pd][2jh+ph][2jw+pw] (0 when a tap index is 5+)
    for (int idx = threadIdx.x; idx < 36 * 64 * 8; idx += kFpThreads) {
        const int c = idx & 7, co = (idx >> 3) & 63, kc = idx >> 9;
        const int jw = kc & 3, jh = (kc >> 2) % 3, jd = (kc >> 2) / 3;
        const int kd = 2 * jd + (c >> 2), kh = 2 * jh + ((c >> 1) & 1), kw = 2 * jw + (c & 1);
        const float v = (kd < 5 && kh < 5 && kw < 5) ? w[co * 125 + (kd * 5 + kh) * 5 + kw] : 0.f;
        reinterpret_cast<__nv_bfloat16*>(b_tile)[idx] = __float2bfloat16(v);
    }
    for (int i = threadIdx.x; i < kARing * kABytes / 16; i += kFpThreads) reinterpret_cast<uint4*>(a_bufs)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x < 64) sbias[threadIdx.x] = bias != nullptr ? bias[threadIdx.x] : 0.f;
    tc::fence_proxy_async_smem();
    if (threadIdx.x == 0) {
        if (pw.yw == nullptr) tc::tma_prefetch_desc(&tm_y);
        for (int i = 0; i < kARing; ++i) { tc::mbar_init(&a_full[i], 4); tc::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&t_full[i], 1); tc::mbar_init(&t_empty[i], 4); }
        for (int i = 0; i < kRawRing; ++i) { tc::mbar_init(&raw_full[i], 1); tc::mbar_init(&raw_empty[i], 4); }
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc<128>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp < 8) {
        staging_loop(x, g, a_bufs, raws, a_full, a_empty, raw_full, raw_empty, kLineBytes, kABytes);
    } else if (warp == 17) {
        if (lane == 0) raw_producer_loop(x, g, raws, raw_full, raw_empty);
    } else if (warp == 8) {
        const bool leader = tc::elect_one();      // the whole warp walks the schedule, one lane issues
        {
            constexpr uint32_t idesc = tc::make_idesc_bf16(128, 64, 0, 0);
            // no-swizzle K-major descriptors as (low word, constant high word): SBO = 128 B for both operands,
            // LBO = 16 B (A: the two K chunks are consecutive positions) / 1024 B (B)
            constexpr uint32_t kHi = ((128u >> 4) & 0x3FFF) | (1u << 14);
            const uint32_t a_lo_base = tc::desc_lo_sw128(tc::smem_u32(a_bufs), 16);
            const uint32_t b_lo = tc::desc_lo_sw128(tc::smem_u32(b_tile), 1024);
            if ((int)blockIdx.x < g.total_tiles) {
                tc::mbar_wait(&a_full[0], 0);
                tc::mbar_wait(&t_empty[0], 1);
                tc::tc_fence_after();
            }
            for (int i = 0;; ++i) {
                const int tile = blockIdx.x + i * gridDim.x;
                if (tile >= g.total_tiles) break;
                const int b = i & 1, buf = i % kARing;
                const uint32_t a_lo = a_lo_base + buf * (kABytes >> 4);
                const uint32_t d_tmem = tmem_base + b * 64;
                if (leader) {
#pragma unroll
                    for (int t = 0; t < 5; ++t) {            // (jd, jh)
                        const int jd = t / 3, jh = t % 3;
#pragma unroll
                        for (int pr = 0; pr < 2; ++pr)       // K = 16: taps (jw, jw+1) = 2 consecutive positions
                            tc::umma_bf16_lohi(d_tmem, a_lo + (((jd * 4 + jh) * kLineBytes + pr * 32) >> 4), b_lo + (t * 4 + pr * 2) * 64, kHi, idesc, (t | pr) != 0);
                    }
                }
                if (tile + (int)gridDim.x < g.total_tiles) {  // next tile's lines / accumulator, checked in the middle of this tile's MMAs
                    const int i1 = i + 1;
                    tc::mbar_wait(&a_full[i1 % kARing], (i1 / kARing) & 1);
                    tc::mbar_wait(&t_empty[i1 & 1], (((uint32_t)i1 >> 1) & 1) ^ 1);
                    tc::tc_fence_after();
                }
                if (leader) {
#pragma unroll
                    for (int t = 5; t < 9; ++t) {
                        const int jd = t / 3, jh = t % 3;
#pragma unroll
                        for (int pr = 0; pr < 2; ++pr)
                            tc::umma_bf16_lohi(d_tmem, a_lo + (((jd * 4 + jh) * kLineBytes + pr * 32) >> 4), b_lo + (t * 4 + pr * 2) * 64, kHi, idesc, 1);
                    }
                    tc::umma_commit(&a_empty[buf]);
                    tc::umma_commit(&t_full[b]);
                }
            }
        }
    } else {
        // ===== two epilogue groups; group e handles tiles i with (i & 1) == e: accumulator e, staging buffer e =====
        const int e = (warp - 9) >> 2;
        const int q = warp & 3;                          // TMEM lane quarter of this warp
        const int row = q * 32 + lane;                   // row r = line*64 + w
        const int et = (threadIdx.x - 9 * 32) & 127;
        const int c8 = et & 7, part = et >> 3;           // stats: 16-byte channel chunk, row residue mod 16
        uint8_t* stg = staging + e * kT16;
        float s1[8], s2[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) s1[j] = s2[j] = 0.f;
        uint32_t sgn[4] = {0u, 0u, 0u, 0u};             // pooled epilogue: sign-bit flips of this thread's 8 channels (gamma < 0: minimum wins)
        if (pw.yw != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (pw.gamma[c8 * 8 + j] < 0.f) sgn[j >> 1] |= (j & 1) ? 0x80000000u : 0x00008000u;
        }
        TileIter t(g, blockIdx.x + e * gridDim.x, 2 * gridDim.x);
        for (int i = e; t.valid(g); i += 2, t.next(g)) {
            const uint32_t u = i >> 1;
            const int n = t.n, d = t.d, h0 = t.h0(), lines = t.lines(g);
            if (et == 0) tc::tma_store_wait_read();      // this group's previous store has drained the staging buffer
            named_bar(3 + e, 128);
            tc::mbar_wait(&t_full[e], u & 1);
            tc::tc_fence_after();
            {
                float v[64];
                tc::tmem_ld64(tmem_base + e * 64 + ((uint32_t)(q * 32) << 16), v);
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&t_empty[e]);   // accumulator is in registers: the next MMA may start
                uint8_t* blk = stg + row * 128;
#pragma unroll
                for (int c = 0; c < 64; c += 8) {
                    const float4 b0 = *reinterpret_cast<const float4*>(sbias + c), b1 = *reinterpret_cast<const float4*>(sbias + c + 4);
                    uint4 o;
                    __nv_bfloat162* ob = reinterpret_cast<__nv_bfloat162*>(&o);
                    ob[0] = __floats2bfloat162_rn(v[c] + b0.x, v[c + 1] + b0.y);
                    ob[1] = __floats2bfloat162_rn(v[c + 2] + b0.z, v[c + 3] + b0.w);
                    ob[2] = __floats2bfloat162_rn(v[c + 4] + b1.x, v[c + 5] + b1.y);
                    ob[3] = __floats2bfloat162_rn(v[c + 6] + b1.z, v[c + 7] + b1.w);
                    *reinterpret_cast<uint4*>(blk + (((c >> 3) ^ (row & 7)) << 4)) = o;
                }
            }
            tc::fence_proxy_async_smem();
            named_bar(3 + e, 128);
            if (pw.yw == nullptr) {
                if (et == 0) {
                    tc::tma_store_5d(&tm_y, stg, 0, 0, h0, d, n);
                    if (lines == 2) tc::tma_store_5d(&tm_y, stg + 64 * 128, 0, 0, h0 + 1, d, n);
                    tc::tma_store_commit();
                }
            } else {
                // item = (line, window along W, 16-byte channel chunk); 128 threads step by 128 items, so a thread keeps its chunk c8
                const int items = lines * pw.Wp * 8;
                for (int item = et; item < items; item += 128) {
                    const int q = item >> 3;
                    const int ln = q / pw.Wp, wq = q - ln * pw.Wp;
                    const int r0 = ln * 64 + 3 * wq;
                    uint4 v[3];
#pragma unroll
                    for (int k = 0; k < 3; ++k) v[k] = *reinterpret_cast<const uint4*>(stg + (r0 + k) * 128 + ((c8 ^ ((r0 + k) & 7)) << 4));
                    const uint32_t* v0 = reinterpret_cast<const uint32_t*>(&v[0]);
                    const uint32_t* v1 = reinterpret_cast<const uint32_t*>(&v[1]);
                    const uint32_t* v2 = reinterpret_cast<const uint32_t*>(&v[2]);
                    uint4 o;
                    uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
                    uint32_t code = 0u;
#pragma unroll
                    for (int wd = 0; wd < 4; ++wd) {
                        const uint32_t x0 = v0[wd] ^ sgn[wd], x1 = v1[wd] ^ sgn[wd], x2 = v2[wd] ^ sgn[wd];
                        const __nv_bfloat162 b0 = *reinterpret_cast<const __nv_bfloat162*>(&x0), b1 = *reinterpret_cast<const __nv_bfloat162*>(&x1),
                                             b2 = *reinterpret_cast<const __nv_bfloat162*>(&x2);
                        const uint32_t g1 = __hgt2_mask(b1, b0);                       // strictly greater: the first one wins ties
                        const uint32_t g2 = __hgt2_mask(b2, __hmax2(b0, b1));
                        const uint32_t w01 = (v1[wd] & g1) | (v0[wd] & ~g1);
                        ow[wd] = (v2[wd] & g2) | (w01 & ~g2);
                        const uint32_t cc = (g1 & ~g2 & 0x00010001u) | (g2 & 0x00020002u);   // 2-bit column per half
                        code |= ((cc & 3u) | ((cc >> 14) & 0xCu)) << (4 * wd);
                    }
                    const size_t pos = (((size_t)n * g.Do + d) * g.Ho + (h0 + ln)) * pw.Wp + wq;
                    *reinterpret_cast<uint4*>(pw.yw + pos * 64 + c8 * 8) = o;
                    pw.aw[pos * 8 + c8] = (unsigned short)code;
                }
            }
            if (stats_partials != nullptr) {
                // 16-byte shared loads: thread (c8, part) sums 8 channels over rows w = part, part+16, ... of each line
                for (int ln = 0; ln < lines; ++ln) {
#pragma unroll 2
                    for (int wq = part; wq < g.Wo; wq += 16) {
                        const int r = ln * 64 + wq;
                        const uint4 pk = *reinterpret_cast<const uint4*>(stg + r * 128 + ((c8 ^ (r & 7)) << 4));
                        const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 f = __bfloat1622float2(pb[j]);
                            s1[2 * j] += f.x; s1[2 * j + 1] += f.y;
                            s2[2 * j] = fmaf(f.x, f.x, s2[2 * j]); s2[2 * j + 1] = fmaf(f.y, f.y, s2[2 * j + 1]);
                        }
                    }
                }
            }
        }
        if (et == 0) tc::tma_store_wait_read();
        if (stats_partials != nullptr) {
            float* rd = red + (e * 16 + part) * 128;
#pragma unroll
            for (int j = 0; j < 8; ++j) { rd[c8 * 8 + j] = s1[j]; rd[64 + c8 * 8 + j] = s2[j]; }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (stats_partials != nullptr && threadIdx.x < 128) {      // fixed-order reduction over 2 groups x 16 parts
        float acc = 0.f;
        for (int p = 0; p < 32; ++p) acc += red[p * 128 + threadIdx.x];
        stats_partials[(size_t)blockIdx.x * 128 + threadIdx.x] = acc;
    }
    if (warp == 8) {
        tc::tc_fence_after();
        tc::tmem_dealloc<128>(tmem_base);
    }
}

// ======================================================================================= wgrad
// D_jw[L*8 + c][co] (jw = 0..2, three accumulators of 64 columns) += line_L[pos + jw][c] * dY[pos][co]
// A = the s2d lines as an MN-major operand: M chunk = one line L = (jd, y' line) (stride 1 KiB), K = positions (16 B apart);
// line chunk (jd, hl) is the tap (jd, jh = hl) because rows 64..127 of the K range run into the next line, exactly like fprop.
// kFused: `tm_dy` maps the conv OUTPUT y instead of its gradient, and four more warps turn every y tile into the dY tile in place
// in shared memory before the MMAs read it:  dY = cA*y + cB + (this position won its pooling window ? da*scale : 0)  -- the
// BatchNorm backward (batch statistics) + ReLU gate + MaxPool(3,3) scatter of the first block, which otherwise costs a kernel of
// its own that reads y and writes dY (2 x 506 MB per step at batch 16) for this kernel to read back.
struct FusedBwd {
    const unsigned char* amax;       // [N][Dp][Hp][Wp][64] winning tap of every pooling window (255 = gated by ReLU)
    const __nv_bfloat16* da;         // [N][Dp][Hp][Wp][64] gradient of the block output
    const float *scale, *mean, *invstd, *c12;      // BatchNorm forward scale = gamma*invstd, batch mean / invstd, backward means (c1 | c2)
    int Dp, Hp, Wp;
};
constexpr int kFusedThreads = kThreads + 4 * 32;     // warps 13-16: y -> dY transform

template <bool kFused>
__global__ void __launch_bounds__(kFused ? kFusedThreads : kThreads, 1) wgrad_kernel(const uint8_t* __restrict__ x, const __grid_constant__ CUtensorMap tm_dy,
                                                            float* __restrict__ partials /* [grid][3][128][64] */, const Geom g, const FusedBwd fb) {
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint8_t* dy_tiles = smem;                              // [kDyRing][128 pos][64 co] swizzled; K row = line*65 + w, the other rows stay zero
    uint8_t* a_bufs = dy_tiles + kDyRing * kT16;           // [kARing][kWgABytes] s2d lines
    uint64_t* bars = reinterpret_cast<uint64_t*>(a_bufs + kARing * kWgABytes);
    uint64_t *a_full = bars, *a_empty = bars + kARing, *t_full = bars + 2 * kARing, *dy_full = t_full + 1, *dy_empty = dy_full + kDyRing;
    uint64_t *raw_full = dy_empty + kDyRing, *raw_empty = raw_full + kRawRing;
    uint64_t* dy_ready = raw_empty + kRawRing;             // [kDyRing] fused: the transform warps have turned the y tile into dY
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(dy_ready + kDyRing);
    float* bnc = reinterpret_cast<float*>(tmem_ptr + 4);   // [3][64] fused: cA, cB, scale per channel
    uint8_t* raws = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(bnc + 192) + 15) & ~(uintptr_t)15);   // 16-byte vector stores
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler: no per-MMA lane-election loops

    for (int i = threadIdx.x; i < (kDyRing * kT16 + kARing * kWgABytes) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    if (kFused && threadIdx.x < 64) {
        // dy = scale*dyhat + A*y + B   with  A = -scale*invstd*c2,  B = scale*(mean*invstd*c2 - c1)   (same arithmetic as bn_pool.cu)
        const int c = threadIdx.x;
        const float sc = fb.scale[c], is = fb.invstd[c], mu = fb.mean[c], k1 = fb.c12[c], k2 = fb.c12[64 + c];
        bnc[c] = -sc * is * k2;
        bnc[64 + c] = sc * (mu * is * k2 - k1);
        bnc[128 + c] = sc;
    }
    tc::fence_proxy_async_smem();
    if (threadIdx.x == 0) {
        tc::tma_prefetch_desc(&tm_dy);
        for (int i = 0; i < kARing; ++i) { tc::mbar_init(&a_full[i], 4); tc::mbar_init(&a_empty[i], 1); }
        for (int i = 0; i < kDyRing; ++i) { tc::mbar_init(&dy_full[i], 1); tc::mbar_init(&dy_empty[i], 1); tc::mbar_init(&dy_ready[i], 4); }
        for (int i = 0; i < kRawRing; ++i) { tc::mbar_init(&raw_full[i], 1); tc::mbar_init(&raw_empty[i], 4); }
        tc::mbar_init(t_full, 1);
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc<256>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp < 8) {
        staging_loop(x, g, a_bufs, raws, a_full, a_empty, raw_full, raw_empty, kWgLineBytes, kWgABytes);
    } else if (warp == 8) {
        const bool leader = tc::elect_one();
        {
            constexpr uint32_t idesc = tc::make_idesc_bf16(128, 64, 1, 1);
            // descriptors as (low word, constant high word): one 32-bit add per MMA, no 64-bit rebuild
            // A = lines, MN-major no swizzle: M chunks (lines) kWgLineBytes apart (SBO), K groups (8 positions) 128 B apart (LBO)
            // B = dY, MN-major SW128 (N = 64 co = one chunk; 8-position groups 1 KiB apart)
            constexpr uint32_t a_hi = ((kWgLineBytes >> 4) & 0x3FFF) | (1u << 14);
            constexpr uint32_t b_hi = tc::desc_hi_sw128(1024);
            const uint32_t a_lo_base = tc::desc_lo_sw128(tc::smem_u32(a_bufs), 128);
            const uint32_t b_lo_base = tc::desc_lo_sw128(tc::smem_u32(dy_tiles), kT16);
            const bool any = (int)blockIdx.x < g.total_tiles;
            uint64_t* dy_go = kFused ? dy_ready : dy_full;     // what the MMAs wait for: the transformed tile, or the TMA load itself
            if (any) {
                tc::mbar_wait(&a_full[0], 0);
                tc::mbar_wait(&dy_go[0], 0);
                tc::tc_fence_after();
            }
            int i = 0;
            for (;; ++i) {
                const int tile = blockIdx.x + i * gridDim.x;
                if (tile >= g.total_tiles) break;
                const int b = i % kARing;
                const int slot = i % kDyRing;
                const uint32_t a_lo = a_lo_base + b * (kWgABytes >> 4);
                const uint32_t b_lo = b_lo_base + slot * (kT16 >> 4);
                const uint32_t acc0 = i != 0;
                if (leader) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)            // 16 positions per MMA; +256 B (A) / +2048 B (B) per step
#pragma unroll
                        for (int jw = 0; jw < 3; ++jw)
                            tc::umma_bf16_lohi2(tmem_base + jw * 64, a_lo + jw + ks * 16, a_hi, b_lo + ks * 128, b_hi, idesc, ks == 0 ? acc0 : 1u);
                }
                if (tile + (int)gridDim.x < g.total_tiles) {   // next tile's operands, checked in the middle of this tile's MMAs
                    const int i1 = i + 1;
                    tc::mbar_wait(&a_full[i1 % kARing], (i1 / kARing) & 1);
                    tc::mbar_wait(&dy_go[i1 % kDyRing], (i1 / kDyRing) & 1);
                    tc::tc_fence_after();
                }
                if (leader) {
#pragma unroll
                    for (int ks = 4; ks < 8; ++ks)
#pragma unroll
                        for (int jw = 0; jw < 3; ++jw)
                            tc::umma_bf16_lohi2(tmem_base + jw * 64, a_lo + jw + ks * 16, a_hi, b_lo + ks * 128, b_hi, idesc, 1u);
                    tc::umma_commit(&a_empty[b]);
                    tc::umma_commit(&dy_empty[slot]);
                }
            }
            if (leader) tc::umma_commit(t_full);
        }
    } else if (kFused && warp >= 13) {
        // ===== y -> dY in place (128 threads).  A thread owns one 16-byte chunk column (8 channels) and up to three W-triplets of the
        // tile: the 3 positions of a triplet lie in ONE pooling window, so they share one (amax, da) load -- and those loads depend
        // only on the tile's coordinates, so they are issued BEFORE waiting for the y tile (their latency hides behind the TMA load)
        const int tt = threadIdx.x - 13 * 32;
        const int cc = tt & 7, tg = tt >> 3;
        const int trips = (g.Wo + 2) / 3;                      // triplets per output line
        float cA[8], cB[8], sc8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { cA[i] = bnc[cc * 8 + i]; cB[i] = bnc[64 + cc * 8 + i]; sc8[i] = bnc[128 + cc * 8 + i]; }
        // (amax, da) of a tile's triplets: issued one tile AHEAD (their L2 / HBM latency is about one tile period)
        auto fetch = [&](const TileIter& ti, uint2* am, uint4* dv, int* tap_h) {
            const int n = ti.n, d = ti.d, h0 = ti.h0();
            const int dw = d / 3, tap_d = (d - dw * 3) * 9;
            const bool d_ok = dw < fb.Dp;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int t3 = tg + 16 * k;
                const int l = t3 >= trips ? 1 : 0, ww = t3 - l * trips;
                const int h = h0 + l, hw = h / 3;
                tap_h[k] = tap_d + (h - hw * 3) * 3;
                am[k] = make_uint2(0xffffffffu, 0xffffffffu);              // 255 everywhere: nobody wins
                dv[k] = make_uint4(0u, 0u, 0u, 0u);
                if (t3 < 2 * trips && d_ok && h < g.Ho && hw < fb.Hp && ww < fb.Wp) {
                    const int64_t po = ((((int64_t)n * fb.Dp + dw) * fb.Hp + hw) * fb.Wp + ww) * 64 + cc * 8;
                    am[k] = *reinterpret_cast<const uint2*>(fb.amax + po);
                    dv[k] = *reinterpret_cast<const uint4*>(fb.da + po);
                }
            }
        };
        TileIter t(g, blockIdx.x, gridDim.x);
        uint2 am[3], am_n[3];
        uint4 dv[3], dv_n[3];
        int tap_h[3], tap_n[3];
        if (t.valid(g)) fetch(t, am, dv, tap_h);
        for (int i = 0; t.valid(g); ++i) {
            const int slot = i % kDyRing;
            const int h0 = t.h0();
            TileIter tn = t;
            tn.next(g);
            if (tn.valid(g)) fetch(tn, am_n, dv_n, tap_n);
            tc::mbar_wait(&dy_full[slot], (i / kDyRing) & 1);
            uint8_t* tile = dy_tiles + slot * kT16;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int t3 = tg + 16 * k;
                if (t3 >= 2 * trips) break;
                const int l = t3 >= trips ? 1 : 0, ww = t3 - l * trips;
                if (h0 + l >= g.Ho) break;                       // a line past the volume: the load zero-filled it and its dY stays zero
                float dz[8];
                const __nv_bfloat162* dh = reinterpret_cast<const __nv_bfloat162*>(&dv[k]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float2 f = __bfloat1622float2(dh[q]);
                    dz[2 * q] = f.x * sc8[2 * q];
                    dz[2 * q + 1] = f.y * sc8[2 * q + 1];
                }
#pragma unroll
                for (int e = 0; e < 3; ++e) {
                    const int w = 3 * ww + e;
                    if (w >= g.Wo) break;
                    const unsigned tap = (unsigned)(tap_h[k] + e);
                    const int row = l * kWgPitch + w;
                    uint4* p = reinterpret_cast<uint4*>(tile + row * 128 + ((cc ^ (row & 7)) << 4));
                    const uint4 u = *p;
                    const __nv_bfloat162* yh = reinterpret_cast<const __nv_bfloat162*>(&u);
                    uint4 o;
                    __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float2 f = __bfloat1622float2(yh[q]);
                        const unsigned word = q < 2 ? am[k].x : am[k].y;
                        const unsigned b0 = (word >> (16 * (q & 1))) & 255u, b1 = (word >> (16 * (q & 1) + 8)) & 255u;
                        oh[q] = __floats2bfloat162_rn(fmaf(cA[2 * q], f.x, cB[2 * q]) + (b0 == tap ? dz[2 * q] : 0.f),
                                                      fmaf(cA[2 * q + 1], f.y, cB[2 * q + 1]) + (b1 == tap ? dz[2 * q + 1] : 0.f));
                    }
                    *p = o;
                }
            }
            tc::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&dy_ready[slot]);
#pragma unroll
            for (int k = 0; k < 3; ++k) { am[k] = am_n[k]; dv[k] = dv_n[k]; tap_h[k] = tap_n[k]; }
            t = tn;
        }
    } else {
        if (warp == 9 && lane == 0) {
            TileIter t(g, blockIdx.x, gridDim.x);
            for (int i = 0; t.valid(g); ++i, t.next(g)) {
                const int slot = i % kDyRing;
                const int n = t.n, d = t.d, h0 = t.h0();
                tc::mbar_wait(&dy_empty[slot], ((i / kDyRing) & 1) ^ 1);     // dY runs kDyRing tiles ahead of the MMAs
                // one box per output line (rows 0..Wo-1 and 64..64+Wo-1); a line past Ho is zero-filled by the hardware
                if (g.dbg & 4) { tc::mbar_arrive(&dy_full[slot]); continue; }
                tc::mbar_expect_tx(&dy_full[slot], (uint32_t)(2 * g.Wo * 128));
                tc::tma_load_5d(&tm_dy, &dy_full[slot], dy_tiles + slot * kT16, 0, 0, h0, d, n);
                tc::tma_load_5d(&tm_dy, &dy_full[slot], dy_tiles + slot * kT16 + kWgPitch * 128, 0, 0, h0 + 1, d, n);   // K rows 65.. = second line
            }
        }
        if (warp == 10 && lane == 0) raw_producer_loop(x, g, raws, raw_full, raw_empty);
        __syncwarp();
        const int q = warp & 3;
        const int mrow = q * 32 + lane;                   // D row = line*8 + channel
        tc::mbar_wait(t_full, 0);
        tc::tc_fence_after();
#pragma unroll 1
        for (int jw = 0; jw < 3; ++jw) {
            float* dst = partials + (((size_t)blockIdx.x * 3 + jw) * 128 + mrow) * 64;
#pragma unroll 1
            for (int c = 0; c < 64; c += 16) {
                float v[16];
                tc::tmem_ld16(tmem_base + jw * 64 + c + ((uint32_t)(q * 32) << 16), v);
#pragma unroll
                for (int j = 0; j < 16; j += 4) st4(dst + c + j, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc::tc_fence_after();
        tc::tmem_dealloc<256>(tmem_base);
    }
}

// dW[co][kd][kh][kw] = sum_cta partials[cta][jw][(jd*4+jh)*8 + c][co]   with kd = 2jd+pd, kh = 2jh+ph, kw = 2jw+pw, c = pd*4+ph*2+pw.
// Threads run over the partial layout (co fastest) so the reads of every CTA's slab are coalesced; the 8000 results scatter.
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partials, int n_cta, float* __restrict__ grad, float* __restrict__ grad_bias,
                                                           const float* __restrict__ w, float* __restrict__ scores, float sal_scale) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 64 && grad_bias != nullptr) grad_bias[i] = 0.f;   // BatchNorm follows this conv: the bias gradient is identically zero
    if (i >= 3 * 128 * 64) return;
    const int co = i & 63, row = (i >> 6) & 127, jw = i >> 13;
    const int L = row >> 3, c = row & 7;
    const int jd = L >> 2, jh = L & 3;
    const int kd = 2 * jd + (c >> 2), kh = 2 * jh + ((c >> 1) & 1), kw = 2 * jw + (c & 1);
    if (jd > 2 || jh > 2 || kd > 4 || kh > 4 || kw > 4) return;
    float acc = 0.f;
    for (int q = 0; q < n_cta; ++q) acc += partials[(size_t)q * (3 * 128 * 64) + i];
    const int o = co * 125 + (kd * 5 + kh) * 5 + kw;
    grad[o] = acc;
    if (scores != nullptr) scores[o] = fmaf(sal_scale, fabsf(acc * w[o]), scores[o]);      // SNIP |dL/dw * w| straight from the reduction
}

static int make_geom(Geom& g, int N, int Di, int Hi, int Wi, float scale) {
    g.N = N; g.Di = Di; g.Hi = Hi; g.Wi = Wi; g.scale = scale;
    g.Do = (Di - 5) / 2 + 1; g.Ho = (Hi - 5) / 2 + 1; g.Wo = (Wi - 5) / 2 + 1;
    if (g.Wo < 1 || g.Wo > 61 || g.Do < 1 || g.Ho < 1) { nidt_set_error("conv_in_s2d: output width %d unsupported (1..61)", g.Wo); return 1; }
    g.h_tiles = (g.Ho + 1) / 2;
    g.total_tiles = N * g.Do * g.h_tiles;
    g.raw_pitch = ((kRows * Wi + 15 + 15) / 16 + 1) * 16;
    g.in_bytes = (long long)N * Di * Hi * Wi;
    { const char* e = getenv("NIDT_S2D_DEBUG"); g.dbg = e ? atoi(e) : 0; }
    return 0;
}
static size_t smem_fprop(const Geom& g) { return 2 * (size_t)kT16 + 36 * 1024 + kARing * (size_t)kABytes + 8 * (2 * kARing + 4 + 2 * kRawRing) + 16 + 256 + 2 * 16 * 2 * 64 * 4 + kRawRing * (size_t)kPlanes * g.raw_pitch + 64; }
static size_t smem_wgrad(const Geom& g) { return kDyRing * (size_t)kT16 + kARing * (size_t)kWgABytes + 8 * (2 * kARing + 1 + 3 * kDyRing + 2 * kRawRing) + 32 + 768 + kRawRing * (size_t)kPlanes * g.raw_pitch + 64; }

}  // namespace s2d
}  // namespace nidt
using namespace nidt;

NIDT_API int nidt_conv_in_s2d_supported(int Di, int Hi, int Wi, int k, int stride, int pad) {
    s2d::Geom g;
    return (k == 5 && stride == 2 && pad == 0 && s2d::make_geom(g, 1, Di, Hi, Wi, 1.f) == 0) ? 1 : 0;
}
NIDT_API int64_t nidt_conv_in_s2d_wgrad_scratch_floats() { return (int64_t)kNumSMs * 3 * 128 * 64; }

static int s2d_fprop_impl(const void* x, const float* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                          float scale, void* stream, const s2d::PoolW& pw);
// same contract as nidt_conv_in_fprop, for k = 5, stride = 2, pad = 0
NIDT_API int nidt_conv_in_s2d_fprop(const void* x, const float* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                                    float scale, void* stream) {
    NIDT_REQUIRE(y != nullptr, "nidt_conv_in_s2d_fprop: y is null");
    return s2d_fprop_impl(x, w, bias, y, stats_partials, N, Di, Hi, Wi, scale, stream, s2d::PoolW{nullptr, nullptr, nullptr, 0});
}
// The same convolution with its output reduced over every 3 consecutive output columns in the epilogue (see s2d::PoolW): yw is bf16
// [N][Do][Ho][Wp][64] with Wp = (Wo - 3) / 3 + 1, aw uint16 [N][Do][Ho][Wp][8] (2 bits per channel: which of the 3 columns), gamma the
// BatchNorm weight that follows (its sign selects maximum / minimum).  The full-resolution output is not written.
NIDT_API int nidt_conv_in_s2d_fprop_poolw(const void* x, const float* w, const float* bias, const float* gamma, void* yw, void* aw,
                                          float* stats_partials, int N, int Di, int Hi, int Wi, float scale, void* stream) {
    NIDT_REQUIRE(yw != nullptr && aw != nullptr && gamma != nullptr, "nidt_conv_in_s2d_fprop_poolw: null argument");
    const int Wo = (Wi - 5) / 2 + 1;
    NIDT_REQUIRE(Wo >= 3, "nidt_conv_in_s2d_fprop_poolw: output width %d < 3", Wo);
    return s2d_fprop_impl(x, w, bias, nullptr, stats_partials, N, Di, Hi, Wi, scale, stream,
                          s2d::PoolW{(__nv_bfloat16*)yw, (unsigned short*)aw, gamma, (Wo - 3) / 3 + 1});
}
static int s2d_fprop_impl(const void* x, const float* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                          float scale, void* stream, const s2d::PoolW& pw) {
    s2d::Geom g;
    if (s2d::make_geom(g, N, Di, Hi, Wi, scale)) return 2;
    CUtensorMap ty;
    memset(&ty, 0, sizeof(ty));
    if (y != nullptr && make_tmap_act5d(&ty, y, N, g.Do, g.Ho, g.Wo, 64, g.Wo, 1, 1, 1, 1)) return 1;
    const size_t smem = s2d::smem_fprop(g);
    NIDT_REQUIRE(smem <= 227 * 1024, "conv_in_s2d_fprop: shared memory %zu B", smem);
    const int grid = g.total_tiles < kNumSMs ? g.total_tiles : kNumSMs;
    cudaStream_t s = (cudaStream_t)stream;
    NIDT_CUDA_CHECK(cudaFuncSetAttribute(s2d::fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    s2d::fprop_kernel<<<grid, s2d::kFpThreads, smem, s>>>((const uint8_t*)x, w, bias, ty, stats_partials, g, pw);
    NIDT_LAUNCH_CHECK();
    if (stats_partials != nullptr && grid < kNumSMs)
        NIDT_CUDA_CHECK(cudaMemsetAsync(stats_partials + (size_t)grid * 128, 0, (size_t)(kNumSMs - grid) * 128 * sizeof(float), s));
    return 0;
}

// scratch: >= nidt_conv_in_s2d_wgrad_scratch_floats() floats
static int s2d_wgrad_impl(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi, int Wi,
                          float scale, void* stream, const float* sal_w, float* sal_scores, float sal_scale, const s2d::FusedBwd* fb = nullptr);
// BatchNorm-backward + ReLU gate + MaxPool(3,3) scatter fused into the weight gradient of the first layer: takes the conv OUTPUT y
// (bf16 [N][Do][Ho][Wo][64]), the winning taps `amax` and the pooled gradient `da` instead of dY; scale / mean / invstd are the
// BatchNorm forward's per-channel values, c12 the two backward means (nidt_bn_relu_pool3_bwd_stats).  sal_w / sal_scores may be
// NULL (else the SNIP epilogue runs in the split reduction).  Same result as nidt_bn_relu_pool_bwd + nidt_conv_in_s2d_wgrad without
// dY ever touching HBM.  ref: autograd of salient_models.py:147-150.
NIDT_API int nidt_conv_in_s2d_wgrad_fused(const void* x, const void* y, const void* amax, const void* da, const float* bn_scale,
                                          const float* bn_mean, const float* bn_invstd, const float* c12, float* grad, float* grad_bias,
                                          float* scratch, int N, int Di, int Hi, int Wi, float scale, const float* sal_w, float* sal_scores,
                                          float sal_scale, void* stream) {
    NIDT_REQUIRE(amax != nullptr && da != nullptr && bn_scale != nullptr && bn_mean != nullptr && bn_invstd != nullptr && c12 != nullptr,
                 "conv_in_s2d_wgrad_fused: null argument");
    s2d::FusedBwd fb;
    fb.amax = (const unsigned char*)amax; fb.da = (const __nv_bfloat16*)da;
    fb.scale = bn_scale; fb.mean = bn_mean; fb.invstd = bn_invstd; fb.c12 = c12;
    fb.Dp = (((Di - 5) / 2 + 1) - 3) / 3 + 1; fb.Hp = (((Hi - 5) / 2 + 1) - 3) / 3 + 1; fb.Wp = (((Wi - 5) / 2 + 1) - 3) / 3 + 1;
    return s2d_wgrad_impl(x, y, grad, grad_bias, scratch, N, Di, Hi, Wi, scale, stream, sal_w, sal_scores, sal_scale, &fb);
}
NIDT_API int nidt_conv_in_s2d_wgrad(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi, int Wi,
                                    float scale, void* stream) {
    return s2d_wgrad_impl(x, dy, grad, grad_bias, scratch, N, Di, Hi, Wi, scale, stream, nullptr, nullptr, 0.f);
}
// the same with the SNIP saliency epilogue in the split reduction: scores[co][125] += sal_scale * |dW * w|  (ref: snip.py:60-71)
NIDT_API int nidt_conv_in_s2d_wgrad_saliency(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi,
                                             int Wi, float scale, const float* w, float* scores, float sal_scale, void* stream) {
    NIDT_REQUIRE(w != nullptr && scores != nullptr, "conv_in_s2d_wgrad_saliency: w / scores is null");
    return s2d_wgrad_impl(x, dy, grad, grad_bias, scratch, N, Di, Hi, Wi, scale, stream, w, scores, sal_scale);
}
static int s2d_wgrad_impl(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi, int Wi,
                          float scale, void* stream, const float* sal_w, float* sal_scores, float sal_scale, const s2d::FusedBwd* fb) {
    s2d::Geom g;
    if (s2d::make_geom(g, N, Di, Hi, Wi, scale)) return 2;
    CUtensorMap tdy;
    if (make_tmap_act5d(&tdy, dy, N, g.Do, g.Ho, g.Wo, 64, g.Wo, 1, 1, 1, 1)) return 1;
    const size_t smem = s2d::smem_wgrad(g);
    NIDT_REQUIRE(smem <= 227 * 1024, "conv_in_s2d_wgrad: shared memory %zu B", smem);
    const int grid = g.total_tiles < kNumSMs ? g.total_tiles : kNumSMs;
    cudaStream_t s = (cudaStream_t)stream;
    if (fb != nullptr) {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(s2d::wgrad_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        s2d::wgrad_kernel<true><<<grid, s2d::kFusedThreads, smem, s>>>((const uint8_t*)x, tdy, scratch, g, *fb);
    } else {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(s2d::wgrad_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        s2d::wgrad_kernel<false><<<grid, s2d::kThreads, smem, s>>>((const uint8_t*)x, tdy, scratch, g, s2d::FusedBwd{});
    }
    NIDT_LAUNCH_CHECK();
    s2d::wgrad_reduce_kernel<<<(3 * 128 * 64 + 255) / 256, 256, 0, s>>>(scratch, grid, grad, grad_bias, sal_w, sal_scores, sal_scale);
    NIDT_LAUNCH_CHECK();
    return 0;
}
