// First-layer 3-D convolution (Cin = 1 -> 64 channels) on the tensor cores.
//
// With one input channel the implicit-GEMM K dimension is the filter volume (5^3 = 125 -> 128), so
// there is no channel vector for TMA to fetch: the im2col tile is *built* in shared memory, directly
// in the 128B-swizzled UMMA layout, by two builder warpgroups from an 8-bit input patch that is
// normalised (u8 * scale -> bf16) while it is staged.  The same builder feeds
//   fprop : Y[pos][co]  = A[pos][tap] . W[co][tap]          (A K-major,  M = 128 positions)
//   wgrad : dW[tap][co] = A[pos][tap]^T . dY[pos][co]       (A MN-major, K = 128 positions, dY by TMA)
// so the 8-bit volume is the only form of the input that ever exists in HBM (the reference copies
// fp32 volumes, 4x the bytes, fedml_api/standalone/sailentgrads/my_model_trainer.py:185-199).
// K slot k^3 of the im2col tile is a constant 1, so the bias rides the GEMM in fprop (W[co][k^3] = bias)
// and its gradient (sum of dY) falls out of wgrad as row k^3.  fprop's epilogue stores bf16 NDHWC tiles
// by TMA and accumulates the BatchNorm (sum, sum^2) partials per persistent CTA.
// Reference op: salient_models.py:147 (K1/K2 of SURVEY §2.5, cuDNN fp32 there).
#include "tc_common.cuh"

namespace nidt {

struct InConvGeom {
    int N, Di, Hi, Wi;      // input volume (1 channel, u8)
    int Do, Ho, Wo;         // output extents
    int stride, pad;
    int TH;                 // output rows (h) per tile: TH*Wo <= 128
    int h_tiles;            // ceil(Ho / TH)
    int PH, PW;             // patch rows = (TH-1)*stride + k ; row pitch in elements (multiple of 4)
    int raw_pitch;          // bytes per plane of the raw u8 staging buffer (multiple of 16)
    int total_tiles;        // N * Do * h_tiles
    float scale;            // x = u8 * scale
};

constexpr int kInThreads = 13 * 32;   // warps 0-3 / 4-7: builder groups, warp 8: MMA, warps 9-12: epilogue (+TMA)
constexpr int kT16 = 16384;           // one [128 rows][64 bf16] swizzled sub-tile

__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// ---- staging of the u8 input patch of one tile as normalised bf16: patch[KS][PH][PW]
// pad == 0 (ST != 0 promises it): the PH input rows of a plane are one contiguous byte run.  It is fetched with
// aligned 16-byte loads into REGISTERS one tile ahead (raw_prefetch), parked in a raw u8 smem buffer (raw_store)
// and converted by a pass in which consecutive lanes handle consecutive 4-element groups (conflict-free 8-byte
// shared stores).  Patch rows past the volume are zero.  pad != 0: element-wise gather with bounds checks.
constexpr int kRawVecs = 3;   // 16-byte vectors per builder thread and tile (KS * raw_pitch / 16 <= 128 * kRawVecs)

__device__ __forceinline__ const uint8_t* plane_src(const uint8_t* __restrict__ x, const InConvGeom& g, int n, int z, int y0) {
    return x + (((size_t)n * g.Di + z) * g.Hi + y0) * g.Wi;
}

template <int KS, int ST>
__device__ __forceinline__ void raw_prefetch(const uint8_t* __restrict__ x, const InConvGeom& g, int n, int d, int h0, int gt, uint4 (&q)[kRawVecs]) {
    const int stride = ST ? ST : g.stride;
    const int z0 = d * stride, y0 = h0 * stride;
    const int bytes = min(g.PH, g.Hi - y0) * g.Wi;
    const int nv = g.raw_pitch >> 4;
#pragma unroll
    for (int j = 0; j < kRawVecs; ++j) {
        const int v = gt + j * 128;
        q[j] = make_uint4(0, 0, 0, 0);
        if (v < KS * nv) {
            const int plane = v / nv, vi = v - plane * nv;
            const uint8_t* src = plane_src(x, g, n, z0 + plane, y0);
            const int mis = (int)(reinterpret_cast<uintptr_t>(src) & 15);
            if (vi * 16 - mis < bytes) q[j] = __ldg(reinterpret_cast<const uint4*>(src - mis) + vi);
        }
    }
}

template <int KS>
__device__ __forceinline__ void raw_store(uint8_t* __restrict__ raw, const InConvGeom& g, int gt, const uint4 (&q)[kRawVecs]) {
    const int nv = g.raw_pitch >> 4;
#pragma unroll
    for (int j = 0; j < kRawVecs; ++j) {
        const int v = gt + j * 128;
        if (v < KS * nv) reinterpret_cast<uint4*>(raw)[v] = q[j];     // raw[plane][vi] with pitch nv vectors == linear index v
    }
}

template <int KS, int PHT, int PWT, int ST>
__device__ __forceinline__ void convert_patch(const uint8_t* __restrict__ x, const uint8_t* __restrict__ raw, __nv_bfloat16* __restrict__ patch,
                                              const InConvGeom& g, int n, int d, int h0, int gt) {
    const int PH = PHT ? PHT : g.PH, PW = PWT ? PWT : g.PW;
    const int stride = ST ? ST : g.stride;
    const int z0 = d * stride, y0 = h0 * stride;
    const int rows = min(PH, g.Hi - y0);
    const int QW = PW >> 2;
    for (int item = gt; item < KS * PH * QW; item += 128) {
        const int R = item / QW, q = item - R * QW;
        const int plane = R / PH, hr = R - plane * PH;
        const int mis = (int)(reinterpret_cast<uintptr_t>(plane_src(x, g, n, z0 + plane, y0)) & 15);
        const int a = plane * g.raw_pitch + mis + hr * g.Wi + 4 * q;     // address of the group's first byte in the raw buffer
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(raw + (a & ~3));
        const uint32_t four = __funnelshift_r(rw[0], rw[1], (a & 3) * 8);
        // u8 -> float without the slow I2F path: float(0x4B000000 | b) - 2^23 == b exactly
        const int nvalid = hr < rows ? min(4, max(0, g.Wi - 4 * q)) : 0;
        const uint32_t fm = four & (nvalid >= 4 ? 0xffffffffu : ((1u << (8 * nvalid)) - 1u));
        float f[4];
        f[0] = (__uint_as_float(__byte_perm(fm, 0x4B000000u, 0x7650)) - 8388608.f) * g.scale;
        f[1] = (__uint_as_float(__byte_perm(fm, 0x4B000000u, 0x7651)) - 8388608.f) * g.scale;
        f[2] = (__uint_as_float(__byte_perm(fm, 0x4B000000u, 0x7652)) - 8388608.f) * g.scale;
        f[3] = (__uint_as_float(__byte_perm(fm, 0x4B000000u, 0x7653)) - 8388608.f) * g.scale;
        __nv_bfloat162 p0 = __floats2bfloat162_rn(f[0], f[1]), p1 = __floats2bfloat162_rn(f[2], f[3]);
        uint2 o2;
        o2.x = *reinterpret_cast<uint32_t*>(&p0);
        o2.y = *reinterpret_cast<uint32_t*>(&p1);
        *reinterpret_cast<uint2*>(patch + R * PW + 4 * q) = o2;
    }
}

template <int KS, int PHT, int PWT, int ST>
__device__ __forceinline__ void gather_patch(const uint8_t* __restrict__ x, __nv_bfloat16* __restrict__ patch, const InConvGeom& g, int n, int d, int h0, int gt) {
    const int PH = PHT ? PHT : g.PH, PW = PWT ? PWT : g.PW;
    const int z0 = d * g.stride - g.pad, y0 = h0 * g.stride - g.pad;
    const int total = KS * PH * PW;
    for (int idx = gt; idx < total; idx += 128) {
        const int w = idx % PW;
        const int t = idx / PW;
        const int hr = t % PH, kd = t / PH;
        const int iz = z0 + kd, iy = y0 + hr, ix = w - g.pad;
        float v = 0.f;
        if (iz >= 0 && iz < g.Di && iy >= 0 && iy < g.Hi && ix >= 0 && ix < g.Wi)
            v = (float)x[(((size_t)n * g.Di + iz) * g.Hi + iy) * g.Wi + ix] * g.scale;
        patch[idx] = __float2bfloat16(v);
    }
}

// ---- one im2col row (= one output position, 128 K slots) into the two swizzled sub-tiles of an A tile.
// K slot kTaps holds the constant 1 (bias column); rows >= valid_rows are all zero.
template <int KS, int PHT, int PWT, int ST>
__device__ __forceinline__ void build_row(const __nv_bfloat16* __restrict__ patch, const InConvGeom& g, int valid_rows, uint8_t* __restrict__ a_tile, int r) {
    constexpr int kTaps = KS * KS * KS;
    static_assert(kTaps < 128, "need a spare K slot for the bias column");
    const int PH = PHT ? PHT : g.PH, PW = PWT ? PWT : g.PW;
    const int sw = r & 7;
    uint8_t* row0 = a_tile + r * 128;
    if (r >= valid_rows) {
#pragma unroll
        for (int chunk = 0; chunk < 16; ++chunk)
            *reinterpret_cast<uint4*>(row0 + (chunk >> 3) * kT16 + (((chunk & 7) ^ sw) << 4)) = make_uint4(0, 0, 0, 0);
        return;
    }
    const int stride = ST ? ST : g.stride;
    const int hh = r / g.Wo, w = r - hh * g.Wo;
    const int e_base = (hh * stride) * PW + w * stride;
    if ((stride & 1) == 0) {
        // even stride: every run of KS taps starts on a 4-byte boundary -> paired (32-bit) shared loads
        const uint32_t* base32 = reinterpret_cast<const uint32_t*>(patch) + (e_base >> 1);
        const int pw2 = PW >> 1;
#pragma unroll
        for (int chunk = 0; chunk < 16; ++chunk) {
            uint32_t wv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int q0 = chunk * 8 + 2 * j, q1 = q0 + 1;
                uint32_t s0 = 0, s1 = 0;
                if (q0 < kTaps) s0 = base32[((q0 / (KS * KS)) * PH + (q0 / KS) % KS) * pw2 + (q0 % KS) / 2];
                if (q1 < kTaps) s1 = base32[((q1 / (KS * KS)) * PH + (q1 / KS) % KS) * pw2 + (q1 % KS) / 2];
                const uint32_t sel = (((q0 % KS) & 1) ? 0x32u : 0x10u) | ((((q1 % KS) & 1) ? 0x76u : 0x54u) << 8);
                wv[j] = __byte_perm(s0, s1, sel);
                if (q1 >= kTaps) wv[j] = (wv[j] & 0xffffu) | (q1 == kTaps ? 0x3F800000u : 0u);
                if (q0 >= kTaps) wv[j] = (q0 == kTaps ? 0x3F80u : 0u);
            }
            *reinterpret_cast<uint4*>(row0 + (chunk >> 3) * kT16 + (((chunk & 7) ^ sw) << 4)) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        }
        return;
    }
    const unsigned short* base = reinterpret_cast<const unsigned short*>(patch) + e_base;
#pragma unroll
    for (int chunk = 0; chunk < 16; ++chunk) {
        uint32_t wv[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int e0 = chunk * 8 + 2 * j, e1 = e0 + 1;
            uint32_t lo = e0 == kTaps ? 0x3F80u : 0u, hi = e1 == kTaps ? 0x3F80u : 0u;
            if (e0 < kTaps) lo = base[((e0 / (KS * KS)) * PH + (e0 / KS) % KS) * PW + e0 % KS];
            if (e1 < kTaps) hi = base[((e1 / (KS * KS)) * PH + (e1 / KS) % KS) * PW + e1 % KS];
            wv[j] = lo | (hi << 16);
        }
        *reinterpret_cast<uint4*>(row0 + (chunk >> 3) * kT16 + (((chunk & 7) ^ sw) << 4)) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
    }
}

__device__ __forceinline__ void decode_tile(const InConvGeom& g, int tile, int& n, int& d, int& h0, int& valid_rows) {
    const int ht = tile % g.h_tiles;
    const int t2 = tile / g.h_tiles;
    d = t2 % g.Do;
    n = t2 / g.Do;
    h0 = ht * g.TH;
    const int rows = min(g.TH, g.Ho - h0);
    valid_rows = rows * g.Wo;
}

__device__ __forceinline__ int patch_elems(const InConvGeom& g, int ks) { return (ks * g.PH * g.PW + 7) & ~7; }

// builder warps (0-7): two groups of 128 threads alternate tiles; group `grp` owns A tile / patch / raw buffer `grp`.
// Software pipeline per group: the global loads of the group's NEXT tile are in flight while the current tile is
// converted and built, and the A tile is only waited for (MMA of two tiles ago finished) right before it is rewritten.
template <int KS, int PHT, int PWT, int ST>
__device__ __forceinline__ void builder_loop(const uint8_t* __restrict__ x, const InConvGeom& g, uint8_t* a_tiles, uint8_t* raws, __nv_bfloat16* patches,
                                             uint64_t* a_full, uint64_t* a_empty) {
    const int grp = threadIdx.x >> 7, gt = threadIdx.x & 127;
    __nv_bfloat16* patch = patches + grp * patch_elems(g, KS);
    uint8_t* raw = raws + grp * KS * g.raw_pitch;
    uint8_t* a_tile = a_tiles + grp * 2 * kT16;
    const bool fast = (ST != 0) || (g.pad == 0 && KS * (g.raw_pitch >> 4) <= 128 * kRawVecs);
    uint4 q[kRawVecs];
    int n, d, h0, valid;
    int tile = blockIdx.x + grp * gridDim.x;
    if (fast && tile < g.total_tiles) {
        decode_tile(g, tile, n, d, h0, valid);
        raw_prefetch<KS, ST>(x, g, n, d, h0, gt, q);
    }
    for (int i = grp; tile < g.total_tiles; i += 2) {
        const uint32_t u = i >> 1;
        decode_tile(g, tile, n, d, h0, valid);
        const int next = tile + 2 * gridDim.x;
        if (fast) {
            raw_store<KS>(raw, g, gt, q);
            named_bar(1 + grp, 128);
            if (next < g.total_tiles) {
                int n2, d2, h2, v2;
                decode_tile(g, next, n2, d2, h2, v2);
                raw_prefetch<KS, ST>(x, g, n2, d2, h2, gt, q);
            }
            convert_patch<KS, PHT, PWT, ST>(x, raw, patch, g, n, d, h0, gt);
        } else {
            gather_patch<KS, PHT, PWT, ST>(x, patch, g, n, d, h0, gt);
        }
        named_bar(1 + grp, 128);
        tc::mbar_wait(&a_empty[grp], (u & 1) ^ 1);
        build_row<KS, PHT, PWT, ST>(patch, g, valid, a_tile, gt);
        tc::fence_proxy_async_smem();
        tc::mbar_arrive(&a_full[grp]);
        tile = next;
    }
}

// =========================================================================== fprop
template <int KS, int PHT, int PWT, int ST>
__global__ void __launch_bounds__(kInThreads, 1) conv_in_fprop_kernel(
    const uint8_t* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, const __grid_constant__ CUtensorMap tm_y,
    float* __restrict__ stats_partials, const InConvGeom g) {
    constexpr int kTaps = KS * KS * KS;
    extern __shared__ __align__(1024) uint8_t smem[];   // SWIZZLE_128B tiles need 1024-byte alignment
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint8_t* a_tiles = smem;                       // [2][2*kT16]
    uint8_t* b_tile = smem + 4 * kT16;             // [2 sub][64 rows][128 B]
    uint8_t* staging = b_tile + kT16;              // [2][128][128 B] swizzled, double buffered
    uint64_t* bars = reinterpret_cast<uint64_t*>(staging + 2 * kT16);
    uint64_t *a_full = bars, *a_empty = bars + 2, *t_full = bars + 4, *t_empty = bars + 6;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 8);
    float* red = reinterpret_cast<float*>(tmem_ptr + 4);     // [4][2][64]
    uint8_t* raws = reinterpret_cast<uint8_t*>(red + 512);
    __nv_bfloat16* patches = reinterpret_cast<__nv_bfloat16*>(raws + 2 * KS * g.raw_pitch);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler: no per-MMA lane-election loops
    // weights (+ bias in K slot kTaps) -> bf16 K-major swizzled B tile (rows = co)
    for (int idx = threadIdx.x; idx < 64 * 128; idx += kInThreads) {
        const int co = idx >> 7, kk = idx & 127;
        float v = 0.f;
        if (kk < kTaps) v = w[co * kTaps + kk];
        else if (kk == kTaps && bias != nullptr) v = bias[co];
        const int sub = kk >> 6, kc = kk & 63;
        *reinterpret_cast<__nv_bfloat16*>(b_tile + sub * (kT16 / 2) + co * 128 + ((((kc >> 3) ^ (co & 7))) << 4) + (kc & 7) * 2) = __float2bfloat16(v);
    }
    tc::fence_proxy_async_smem();
    if (threadIdx.x == 0) {
        tc::tma_prefetch_desc(&tm_y);
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(&a_full[i], 128); tc::mbar_init(&a_empty[i], 1);
            tc::mbar_init(&t_full[i], 1);   tc::mbar_init(&t_empty[i], 128);
        }
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc<128>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp < 8) {
        builder_loop<KS, PHT, PWT, ST>(x, g, a_tiles, raws, patches, a_full, a_empty);
    } else if (warp == 8) {
        const bool leader = tc::elect_one();      // the whole warp walks the schedule, one lane issues
        {
            constexpr uint32_t idesc = tc::make_idesc_bf16(128, 64, 0, 0);
            for (int i = 0;; ++i) {
                const int tile = blockIdx.x + i * gridDim.x;
                if (tile >= g.total_tiles) break;
                const int b = i & 1;
                const uint32_t u = i >> 1;
                tc::mbar_wait(&a_full[b], u & 1);
                tc::mbar_wait(&t_empty[b], (u & 1) ^ 1);
                tc::tc_fence_after();
                const uint32_t a_addr = tc::smem_u32(a_tiles + b * 2 * kT16);
                const uint32_t b_addr = tc::smem_u32(b_tile);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint64_t da = tc::make_smem_desc_sw128(a_addr + (k >> 2) * kT16 + (k & 3) * 32, 16, 1024);
                    const uint64_t db = tc::make_smem_desc_sw128(b_addr + (k >> 2) * (kT16 / 2) + (k & 3) * 32, 16, 1024);
                    if (leader) tc::umma_bf16(tmem_base + b * 64, da, db, idesc, k != 0);
                }
                if (leader) {
                    tc::umma_commit(&a_empty[b]);
                    tc::umma_commit(&t_full[b]);
                }
            }
        }
    } else {
        // ===== epilogue =====
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int et = threadIdx.x - 9 * 32;
        const int ch = et & 63, part = et >> 6;      // stats: channel, row residue (0/1)
        float s1 = 0.f, s2 = 0.f;
        for (int i = 0;; ++i) {
            const int tile = blockIdx.x + i * gridDim.x;
            if (tile >= g.total_tiles) break;
            const int b = i & 1;
            const uint32_t u = i >> 1;
            int n, d, h0, valid;
            decode_tile(g, tile, n, d, h0, valid);
            uint8_t* stg = staging + b * kT16;
            // the TMA store issued two tiles ago must have finished reading this staging buffer
            if (et == 0) tc::tma_store_wait_read_n<1>();
            named_bar(3, 128);
            tc::mbar_wait(&t_full[b], u & 1);
            tc::tc_fence_after();
#pragma unroll
            for (int c = 0; c < 64; c += 16) {
                float v[16];
                tc::tmem_ld16(tmem_base + b * 64 + c + ((uint32_t)(q * 32) << 16), v);
                uint4 o[2];
                __nv_bfloat162* ob = reinterpret_cast<__nv_bfloat162*>(o);
#pragma unroll
                for (int j = 0; j < 8; ++j) ob[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                uint8_t* blk = stg + row * 128;
                const int c0 = c >> 3;
                *reinterpret_cast<uint4*>(blk + ((c0 ^ (row & 7)) << 4)) = o[0];
                *reinterpret_cast<uint4*>(blk + (((c0 + 1) ^ (row & 7)) << 4)) = o[1];
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&t_empty[b]);
            tc::fence_proxy_async_smem();
            named_bar(3, 128);
            if (et == 0) {
                tc::tma_store_5d(&tm_y, stg, 0, 0, h0, d, n);
                tc::tma_store_commit();
            }
            if (stats_partials != nullptr) {
                const int chunk = ch >> 3, within = (ch & 7) * 2;
                float t1 = 0.f, t2 = 0.f;
#pragma unroll 4
                for (int r = part; r < valid; r += 2) {
                    const float xv = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(stg + r * 128 + ((chunk ^ (r & 7)) << 4) + within));
                    t1 += xv;
                    t2 = fmaf(xv, xv, t2);
                }
                s1 += t1;
                s2 += t2;
            }
        }
        if (et == 0) tc::tma_store_wait_read();
        if (stats_partials != nullptr) {
            red[(part * 2 + 0) * 64 + ch] = s1;
            red[(part * 2 + 1) * 64 + ch] = s2;
            named_bar(3, 128);
            if (part == 0) {
                stats_partials[((size_t)blockIdx.x * 2 + 0) * 64 + ch] = red[0 * 64 + ch] + red[2 * 64 + ch];
                stats_partials[((size_t)blockIdx.x * 2 + 1) * 64 + ch] = red[1 * 64 + ch] + red[3 * 64 + ch];
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc::tc_fence_after();
        tc::tmem_dealloc<128>(tmem_base);
    }
}

// =========================================================================== wgrad
template <int KS, int PHT, int PWT, int ST>
__global__ void __launch_bounds__(kInThreads, 1) conv_in_wgrad_kernel(
    const uint8_t* __restrict__ x, const __grid_constant__ CUtensorMap tm_dy, float* __restrict__ partials /* [grid][128][64] */, const InConvGeom g) {
    extern __shared__ __align__(1024) uint8_t smem[];   // SWIZZLE_128B tiles need 1024-byte alignment
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint8_t* a_tiles = smem;                   // [2][2*kT16]
    uint8_t* dy_tiles = smem + 4 * kT16;       // [2][kT16]
    uint64_t* bars = reinterpret_cast<uint64_t*>(dy_tiles + 2 * kT16);
    uint64_t *a_full = bars, *a_empty = bars + 2, *dy_full = bars + 4, *t_full = bars + 6;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 8);
    uint8_t* raws = reinterpret_cast<uint8_t*>(tmem_ptr + 4);
    __nv_bfloat16* patches = reinterpret_cast<__nv_bfloat16*>(raws + 2 * KS * g.raw_pitch);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for the compiler: no per-MMA lane-election loops

    for (int i = threadIdx.x; i < 2 * kT16 / 16; i += kInThreads) reinterpret_cast<uint4*>(dy_tiles)[i] = make_uint4(0, 0, 0, 0);
    tc::fence_proxy_async_smem();
    if (threadIdx.x == 0) {
        tc::tma_prefetch_desc(&tm_dy);
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&a_full[i], 128); tc::mbar_init(&a_empty[i], 1); tc::mbar_init(&dy_full[i], 1); }
        tc::mbar_init(t_full, 1);
        tc::mbar_fence_init();
    }
    if (warp == 8) tc::tmem_alloc<64>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp < 8) {
        builder_loop<KS, PHT, PWT, ST>(x, g, a_tiles, raws, patches, a_full, a_empty);
    } else if (warp == 8) {
        const bool leader = tc::elect_one();
        {
            constexpr uint32_t idesc = tc::make_idesc_bf16(128, 64, 1, 1);
            int i = 0;
            for (;; ++i) {
                const int tile = blockIdx.x + i * gridDim.x;
                if (tile >= g.total_tiles) break;
                const int b = i & 1;
                const uint32_t u = i >> 1;
                tc::mbar_wait(&a_full[b], u & 1);
                tc::mbar_wait(&dy_full[b], u & 1);
                tc::tc_fence_after();
                const uint32_t a_addr = tc::smem_u32(a_tiles + b * 2 * kT16);
                const uint32_t b_addr = tc::smem_u32(dy_tiles + b * kT16);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const uint64_t da = tc::make_smem_desc_sw128(a_addr + k * 2048, kT16, 1024);
                    const uint64_t db = tc::make_smem_desc_sw128(b_addr + k * 2048, kT16, 1024);
                    if (leader) tc::umma_bf16(tmem_base, da, db, idesc, (i | k) != 0);
                }
                if (leader) tc::umma_commit(&a_empty[b]);
            }
            if (leader) tc::umma_commit(t_full);
        }
    } else {
        if (warp == 9 && lane == 0) {
            // dY tiles by TMA (box rows = TH*Wo; rows past Ho are zero-filled by the hardware)
            for (int i = 0;; ++i) {
                const int tile = blockIdx.x + i * gridDim.x;
                if (tile >= g.total_tiles) break;
                const int b = i & 1;
                const uint32_t u = i >> 1;
                int n, d, h0, valid;
                decode_tile(g, tile, n, d, h0, valid);
                tc::mbar_wait(&a_empty[b], (u & 1) ^ 1);
                tc::mbar_expect_tx(&dy_full[b], (uint32_t)(g.TH * g.Wo * 128));
                tc::tma_load_5d(&tm_dy, &dy_full[b], dy_tiles + b * kT16, 0, 0, h0, d, n);
            }
        }
        __syncwarp();
        const int q = warp & 3;
        const int tap = q * 32 + lane;
        tc::mbar_wait(t_full, 0);
        tc::tc_fence_after();
        float* dst = partials + ((size_t)blockIdx.x * 128 + tap) * 64;
#pragma unroll 1
        for (int c = 0; c < 64; c += 16) {
            float v[16];
            tc::tmem_ld16(tmem_base + c + ((uint32_t)(q * 32) << 16), v);
#pragma unroll
            for (int j = 0; j < 16; j += 4) st4(dst + c + j, make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        tc::tc_fence_after();
        tc::tmem_dealloc<64>(tmem_base);
    }
}

// dW[co][tap] = sum_cta partials[cta][tap][co] ; dbias[co] = sum_cta partials[cta][taps][co]   (co fastest: coalesced reads)
__global__ void __launch_bounds__(256) conv_in_wgrad_reduce_kernel(const float* __restrict__ partials, int n_cta, int taps, float* __restrict__ grad,
                                                                   float* __restrict__ grad_bias) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 64 * (taps + 1)) return;
    const int co = i & 63, t = i >> 6;
    float acc = 0.f;
    for (int c = 0; c < n_cta; ++c) acc += partials[((size_t)c * 128 + t) * 64 + co];
    if (t < taps) grad[co * taps + t] = acc;
    else if (grad_bias != nullptr) grad_bias[co] = acc;
}

static int make_geom(InConvGeom& g, int N, int Di, int Hi, int Wi, int k, int stride, int pad, float scale) {
    g.N = N; g.Di = Di; g.Hi = Hi; g.Wi = Wi; g.stride = stride; g.pad = pad; g.scale = scale;
    g.Do = (Di + 2 * pad - k) / stride + 1;
    g.Ho = (Hi + 2 * pad - k) / stride + 1;
    g.Wo = (Wi + 2 * pad - k) / stride + 1;
    if (g.Wo < 1 || g.Wo > 128) { nidt_set_error("conv_in: output width %d unsupported (1..128)", g.Wo); return 1; }
    g.TH = 128 / g.Wo;
    if (g.TH > g.Ho) g.TH = g.Ho;
    g.h_tiles = (g.Ho + g.TH - 1) / g.TH;
    g.PH = (g.TH - 1) * stride + k;
    g.PW = (Wi + 2 * pad + 3) & ~3;
    g.raw_pitch = ((g.PH * Wi + 15 + 15) / 16 + 1) * 16;   // run + worst-case misalignment, +1 vector for the funnel read
    g.total_tiles = N * g.Do * g.h_tiles;
    return 0;
}
static size_t in_smem_bytes(const InConvGeom& g, int k) {
    const size_t patch = (((size_t)k * g.PH * g.PW + 7) & ~(size_t)7) * 2;
    return 7 * (size_t)kT16 + 64 + 16 + 2048 + 2 * (size_t)k * g.raw_pitch + 2 * patch + 64;
}

template <int KS, int PHT, int PWT, int ST>
static int launch_fprop(const InConvGeom& g, const void* x, const float* w, const float* bias, const CUtensorMap& ty, float* stats, int grid, size_t smem, cudaStream_t s) {
    NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv_in_fprop_kernel<KS, PHT, PWT, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_in_fprop_kernel<KS, PHT, PWT, ST><<<grid, kInThreads, smem, s>>>((const uint8_t*)x, w, bias, ty, stats, g);
    NIDT_LAUNCH_CHECK();
    return 0;
}
template <int KS, int PHT, int PWT, int ST>
static int launch_wgrad(const InConvGeom& g, const void* x, const CUtensorMap& tdy, float* scratch, int grid, size_t smem, cudaStream_t s) {
    NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv_in_wgrad_kernel<KS, PHT, PWT, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    conv_in_wgrad_kernel<KS, PHT, PWT, ST><<<grid, kInThreads, smem, s>>>((const uint8_t*)x, tdy, scratch, g);
    NIDT_LAUNCH_CHECK();
    return 0;
}

}  // namespace nidt
using namespace nidt;

NIDT_API int nidt_conv_in_grid() { return kNumSMs; }

// x: u8 [N][Di][Hi][Wi]; w: fp32 [64][k^3] (slab layout [co][1][k][k][k]); y: bf16 [N][Do][Ho][Wo][64]
// stats_partials: nullptr or fp32 [nidt_conv_in_grid()][2][64]
NIDT_API int nidt_conv_in_fprop(const void* x, const float* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                                int k, int stride, int pad, float scale, void* stream) {
    NIDT_REQUIRE(k == 5 || k == 3, "conv_in_fprop: filter extent %d unsupported (3 or 5)", k);
    InConvGeom g;
    if (make_geom(g, N, Di, Hi, Wi, k, stride, pad, scale)) return 2;
    CUtensorMap ty;
    if (make_tmap_act5d(&ty, y, N, g.Do, g.Ho, g.Wo, 64, g.Wo, g.TH, 1, 1, 1)) return 1;
    const size_t smem = in_smem_bytes(g, k);
    NIDT_REQUIRE(smem <= 227 * 1024, "conv_in_fprop: patch too large for shared memory (%zu B)", smem);
    const int grid = g.total_tiles < kNumSMs ? g.total_tiles : kNumSMs;
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (k == 5 && g.PH == 7 && g.PW == 124 && stride == 2 && pad == 0) rc = launch_fprop<5, 7, 124, 2>(g, x, w, bias, ty, stats_partials, grid, smem, s);   // ABCD 121^3-class volumes
    else if (k == 5) rc = launch_fprop<5, 0, 0, 0>(g, x, w, bias, ty, stats_partials, grid, smem, s);
    else rc = launch_fprop<3, 0, 0, 0>(g, x, w, bias, ty, stats_partials, grid, smem, s);
    if (rc) return rc;
    if (stats_partials != nullptr && grid < kNumSMs)   // unused partial rows must read as zero
        NIDT_CUDA_CHECK(cudaMemsetAsync(stats_partials + (size_t)grid * 128, 0, (size_t)(kNumSMs - grid) * 128 * sizeof(float), s));
    return 0;
}

// dy: bf16 [N][Do][Ho][Wo][64]; grad: fp32 [64][k^3]; grad_bias: fp32 [64] or null; scratch: >= nidt_conv_in_grid()*128*64 floats
NIDT_API int nidt_conv_in_wgrad(const void* x, const void* dy, float* grad, float* grad_bias, float* scratch, int N, int Di, int Hi, int Wi, int k,
                                int stride, int pad, float scale, void* stream) {
    NIDT_REQUIRE(k == 5 || k == 3, "conv_in_wgrad: filter extent %d unsupported (3 or 5)", k);
    InConvGeom g;
    if (make_geom(g, N, Di, Hi, Wi, k, stride, pad, scale)) return 2;
    CUtensorMap tdy;
    if (make_tmap_act5d(&tdy, dy, N, g.Do, g.Ho, g.Wo, 64, g.Wo, g.TH, 1, 1, 1)) return 1;
    const size_t smem = in_smem_bytes(g, k);
    NIDT_REQUIRE(smem <= 227 * 1024, "conv_in_wgrad: patch too large for shared memory (%zu B)", smem);
    const int grid = g.total_tiles < kNumSMs ? g.total_tiles : kNumSMs;
    cudaStream_t s = (cudaStream_t)stream;
    int rc;
    if (k == 5 && g.PH == 7 && g.PW == 124 && stride == 2 && pad == 0) rc = launch_wgrad<5, 7, 124, 2>(g, x, tdy, scratch, grid, smem, s);
    else if (k == 5) rc = launch_wgrad<5, 0, 0, 0>(g, x, tdy, scratch, grid, smem, s);
    else rc = launch_wgrad<3, 0, 0, 0>(g, x, tdy, scratch, grid, smem, s);
    if (rc) return rc;
    const int taps = k * k * k;
    conv_in_wgrad_reduce_kernel<<<(64 * (taps + 1) + 255) / 256, 256, 0, s>>>(scratch, grid, taps, grad, grad_bias);
    NIDT_LAUNCH_CHECK();
    return 0;
}
