// 3-D convolution as implicit GEMM on the 5th-gen tensor cores (tcgen05.mma, fp32 accumulators in
// TMEM) fed by TMA, for channels-last bf16 activations with Cin % 64 == 0.
//
//   fprop :  Y[m, co]  = sum_{tap, ci} X[m (+) tap, ci] * Wf[co][tap][ci]          (ref K1: nn.Conv3d fwd)
//   dgrad :  the same kernel on dY with tap-flipped, transposed weights and pad' = k-1-pad (ref K2)
//   wgrad :  dW[tap][co][ci] = sum_m dY[m, co] * X[m (+) tap, ci]   (MN-major operands, split-K) (ref K2)
//
// The GEMM M dimension is a *box* of output positions {bw x bh x bd x bn}: one TMA tiled load of the
// 5-D activation tensor per filter tap brings the shifted input box into shared memory, zero-filling
// everything outside the volume -- padding, ragged edge tiles and halos cost no instructions.
// Persistent CTAs (one per SM), warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer,
// warps 2-5 = epilogue (TMEM -> registers -> bf16 -> swizzled smem -> TMA store, plus the
// per-channel (sum, sum^2) partials BatchNorm needs, so the conv output is never re-read for them).
// Reference call sites: fedml_api/model/cv/salient_models.py:152-166 (cuDNN fp32 there).
#include "tc_common.cuh"
#include <mutex>
#include <string.h>

namespace nidt {

// --------------------------------------------------------------------------- tensor maps
PFN_encodeTiled get_encode_tiled() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    });
    return fn;
}

int make_tmap_act5d(CUtensorMap* out, const void* base, int N, int D, int H, int W, int C, int bw, int bh, int bd, int bn, int stride) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) { nidt_set_error("cuTensorMapEncodeTiled is unavailable"); return 1; }
    cuuint64_t gdim[5] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)D, (cuuint64_t)N};
    cuuint64_t gstr[4] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2, (cuuint64_t)D * H * W * C * 2};
    // with an element stride s the box spans (b-1)*s+1 input elements and yields b of them
    cuuint32_t box[5] = {64u, (cuuint32_t)((bw - 1) * stride + 1), (cuuint32_t)((bh - 1) * stride + 1), (cuuint32_t)((bd - 1) * stride + 1),
                         (cuuint32_t)bn};
    cuuint32_t estr[5] = {1u, (cuuint32_t)stride, (cuuint32_t)stride, (cuuint32_t)stride, 1u};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        nidt_set_error("cuTensorMapEncodeTiled(5d) failed: %d (N=%d D=%d H=%d W=%d C=%d box=%d,%d,%d,%d stride=%d)", (int)r, N, D, H, W, C, bw, bh, bd, bn, stride);
        return 1;
    }
    return 0;
}

int make_tmap_mat2d(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int box_rows) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) { nidt_set_error("cuTensorMapEncodeTiled is unavailable"); return 1; }
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { nidt_set_error("cuTensorMapEncodeTiled(2d) failed: %d rows=%lld cols=%lld", (int)r, (long long)rows, (long long)cols); return 1; }
    return 0;
}

struct ConvGeom {
    int N, Do, Ho, Wo, Cout;    // output tensor (GEMM M = N*Do*Ho*Wo, GEMM N = Cout)
    int k, stride, pad;         // filter extent, stride, padding (dgrad passes pad' = k-1-pad)
    int cin_blocks;             // Cin / 64
    int bw, bh, bd, bn;         // M-tile box
    int tw, th, td, tn;         // tiles per dimension
    int n_tiles;                // ceil(Cout / BLOCK_N)
    int box_rows;               // bw*bh*bd*bn (<= 128)
};

constexpr int kConvThreads = 192;
constexpr int kATileBytes = 128 * 128;   // 128 rows x 64 bf16

// ---- epilogue helpers shared by the fprop kernels -------------------------------------------------------------------------
constexpr int kStatsRedBytes = 8 * 128 * 2 * 4;     // per-row-group partial (sum, sum^2) of every channel of a tile

// TMEM accumulator (this thread's row, BLOCK_N fp32 columns) + bias -> bf16 -> the SWIZZLE_128B staging tile of the TMA store
template <int BLOCK_N>
__device__ __forceinline__ void drain_tile(uint32_t tmem_addr, const float* __restrict__ bias_s, uint8_t* __restrict__ staging, int row) {
#pragma unroll 1
    for (int c = 0; c < BLOCK_N; c += 16) {
        float v[16];
        tc::tmem_ld16(tmem_addr + c, v);
        if (bias_s != nullptr) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += bias_s[c + i];
        }
        uint4 o[2];
        __nv_bfloat162* ob = reinterpret_cast<__nv_bfloat162*>(o);
#pragma unroll
        for (int i = 0; i < 8; ++i) ob[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        uint8_t* blk = staging + (c >> 6) * kATileBytes + row * 128;
        const int ch0 = (c & 63) >> 3;
        *reinterpret_cast<uint4*>(blk + (((ch0) ^ (row & 7)) << 4)) = o[0];
        *reinterpret_cast<uint4*>(blk + (((ch0 + 1) ^ (row & 7)) << 4)) = o[1];
    }
}

// Per-channel (sum, sum^2) over the valid rows of the staged bf16 tile.  Thread `et` of the 128 epilogue threads reads 16-byte
// chunks (8 channels of one row): chunk column et % (BLOCK_N/8), rows et / (BLOCK_N/8) + RG*j -- 16 vector loads per thread
// instead of 128 scalar ones -- the row groups are then added up through shared memory (`red`, kStatsRedBytes) in a fixed order.
// Ends with the result of channel `et` (< BLOCK_N) in (s1, s2).  Uses named barrier 1 (128 threads).
template <int BLOCK_N>
__device__ __forceinline__ void tile_stats(const uint8_t* __restrict__ staging, const uint8_t* __restrict__ row_valid, int box_rows,
                                           float* __restrict__ red, int et, float& s1, float& s2) {
    constexpr int kChunks = BLOCK_N / 8;          // 16-byte chunks per row
    constexpr int kRG = 128 / kChunks;            // row groups
    const int cc = et % kChunks, rg = et / kChunks;
    float a1[8], a2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a1[i] = 0.f; a2[i] = 0.f; }
    const uint8_t* blk = staging + (cc >> 3) * kATileBytes;
    const int ch0 = cc & 7;
    for (int r = rg; r < box_rows; r += kRG) {
        if (!row_valid[r]) continue;
        const uint4 u = *reinterpret_cast<const uint4*>(blk + r * 128 + ((ch0 ^ (r & 7)) << 4));
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 f = __bfloat1622float2(h[i]);
            a1[2 * i] += f.x; a2[2 * i] = fmaf(f.x, f.x, a2[2 * i]);
            a1[2 * i + 1] += f.y; a2[2 * i + 1] = fmaf(f.y, f.y, a2[2 * i + 1]);
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        red[(rg * BLOCK_N + cc * 8 + i) * 2 + 0] = a1[i];
        red[(rg * BLOCK_N + cc * 8 + i) * 2 + 1] = a2[i];
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    s1 = 0.f; s2 = 0.f;
    if (et < BLOCK_N) {
#pragma unroll
        for (int g = 0; g < kRG; ++g) { s1 += red[(g * BLOCK_N + et) * 2 + 0]; s2 += red[(g * BLOCK_N + et) * 2 + 1]; }
    }
}

template <int BLOCK_N>
struct FpropCfg {
    static constexpr int kStages = BLOCK_N == 128 ? 5 : 6;
    static constexpr int kBTileBytes = BLOCK_N * 128;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
    static constexpr int kStagingBytes = (BLOCK_N / 64) * kATileBytes;
    static constexpr int kTmemCols = 2 * BLOCK_N;   // double-buffered accumulator (power of two: 128 / 256)
    static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + 256 /*barriers, flags*/ + 128 + 512 /*bias*/ +
                                      kStatsRedBytes;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(kConvThreads, 1) conv3d_fprop_tc_kernel(
    const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_y,
    const ConvGeom g, const float* __restrict__ bias, float* __restrict__ stats_partials, int total_tiles) {
    using Cfg = FpropCfg<BLOCK_N>;
    extern __shared__ __align__(1024) uint8_t smem[];   // SWIZZLE_128B tiles need 1024-byte alignment
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint8_t* stage_base = smem;
    uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
    uint64_t* full = bars;                       // [kStages]
    uint64_t* empty = bars + Cfg::kStages;       // [kStages]
    uint64_t* tmem_full = empty + Cfg::kStages;  // [2]
    uint64_t* tmem_empty = tmem_full + 2;        // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    uint8_t* row_valid = reinterpret_cast<uint8_t*>(tmem_ptr + 2);   // [128]
    float* bias_s = reinterpret_cast<float*>(row_valid + 128);        // [128]: bias of this tile's channels
    float* stat_red = bias_s + 128;                                    // kStatsRedBytes

    // shuffled warp index = provably warp-uniform: the producer / MMA schedules then live in uniform registers and UTMALDG /
    // UTCHMMA are issued without a per-instruction lane-election loop (the whole warp walks the schedule, one lane issues)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int taps = g.k * g.k * g.k;
    const int num_kb = taps * g.cin_blocks;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&tm_x);
        tc::tma_prefetch_desc(&tm_w);
        tc::tma_prefetch_desc(&tm_y);
        for (int i = 0; i < Cfg::kStages; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&tmem_full[i], 1); tc::mbar_init(&tmem_empty[i], 128); }
        tc::mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    auto decode = [&](int tile, int& n_tile, int& w0, int& h0, int& d0, int& n0) {
        n_tile = tile % g.n_tiles;
        int m = tile / g.n_tiles;
        w0 = (m % g.tw) * g.bw; m /= g.tw;
        h0 = (m % g.th) * g.bh; m /= g.th;
        d0 = (m % g.td) * g.bd; m /= g.td;
        n0 = m * g.bn;
    };

    if (warp == 0) {
        // ===== TMA producer (no divisions inside the k loop: a lone warp pays ~5-10 cycles per scalar instruction) =====
        const bool leader = tc::elect_one();
        uint32_t s = 0, ph = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            int n_tile, w0, h0, d0, n0;
            decode(tile, n_tile, w0, h0, d0, n0);
            const int x0 = w0 * g.stride - g.pad, y0 = h0 * g.stride - g.pad, z0 = d0 * g.stride - g.pad;
            int kb = 0;
            for (int kd = 0; kd < g.k; ++kd)
                for (int kh = 0; kh < g.k; ++kh)
                    for (int kw = 0; kw < g.k; ++kw)
                        for (int cb = 0; cb < g.cin_blocks; ++cb, ++kb) {
                            tc::mbar_wait(&empty[s], ph ^ 1);
                            if (leader) {
                                uint8_t* a_dst = stage_base + s * Cfg::kStageBytes;
                                tc::mbar_expect_tx(&full[s], (uint32_t)(g.box_rows * 128 + Cfg::kBTileBytes));
                                tc::tma_load_5d(&tm_x, &full[s], a_dst, cb * 64, x0 + kw, y0 + kh, z0 + kd, n0);
                                tc::tma_load_2d(&tm_w, &full[s], a_dst + kATileBytes, kb * 64, n_tile * BLOCK_N);
                            }
                            if (++s == (uint32_t)Cfg::kStages) { s = 0; ph ^= 1; }
                        }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const bool leader = tc::elect_one();
        constexpr uint32_t idesc = tc::make_idesc_bf16(128, BLOCK_N, 0, 0);
        constexpr uint32_t kDescHi = tc::desc_hi_sw128(1024);
        const uint32_t stage_lo = tc::desc_lo_sw128(tc::smem_u32(stage_base));
        // The wait for the NEXT stage's operands is placed in the middle of the current stage's MMAs: an mbarrier wait +
        // tcgen05.fence in front of a group of MMAs leaves the tensor pipe idle for ~100 cycles (the MMA queue is shallow),
        // in the middle of a group it is free (scripts/bench_native/mma_rate.cu, modes 2 / 3).
        uint32_t s = 0, ph = 0, local = 0;
        if ((int)blockIdx.x < total_tiles) {
            tc::mbar_wait(&full[0], 0);
            tc::tc_fence_after();
        }
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
            const int buf = local & 1;
            tc::mbar_wait(&tmem_empty[buf], ((local >> 1) & 1) ^ 1);
            tc::tc_fence_after();
            const uint32_t d_tmem = tmem_base + buf * BLOCK_N;
            const bool more_tiles = tile + (int)gridDim.x < total_tiles;
            for (int kb = 0; kb < num_kb; ++kb) {
                const uint32_t a_lo = stage_lo + s * (Cfg::kStageBytes >> 4);
                const uint32_t b_lo = a_lo + (kATileBytes >> 4);
                uint32_t s1 = s + 1, ph1 = ph;
                if (s1 == (uint32_t)Cfg::kStages) { s1 = 0; ph1 ^= 1; }
                if (leader) {
                    tc::umma_bf16_lohi(d_tmem, a_lo, b_lo, kDescHi, idesc, kb != 0);
                    tc::umma_bf16_lohi(d_tmem, a_lo + 2, b_lo + 2, kDescHi, idesc, 1);
                }
                if (kb + 1 < num_kb || more_tiles) {
                    tc::mbar_wait(&full[s1], ph1);
                    tc::tc_fence_after();
                }
                if (leader) {
                    tc::umma_bf16_lohi(d_tmem, a_lo + 4, b_lo + 4, kDescHi, idesc, 1);
                    tc::umma_bf16_lohi(d_tmem, a_lo + 6, b_lo + 6, kDescHi, idesc, 1);
                    tc::umma_commit(&empty[s]);
                }
                s = s1; ph = ph1;
            }
            if (leader) tc::umma_commit(&tmem_full[buf]);
        }
    } else {
        // ===== epilogue: warps 2..5; TMEM lane quarter = warp % 4 =====
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int et = threadIdx.x - 64;     // 0..127
        uint32_t local = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++local) {
            int n_tile, w0, h0, d0, n0;
            decode(tile, n_tile, w0, h0, d0, n0);
            const int buf = local & 1;
            const uint32_t aph = (local >> 1) & 1;
            // the previous tile's TMA store must have finished reading the staging buffer
            if (et == 0) tc::tma_store_wait_read();
            const int co0 = n_tile * BLOCK_N;
            if (bias != nullptr && et < BLOCK_N) bias_s[et] = (co0 + et < g.Cout) ? __ldg(bias + co0 + et) : 0.f;
            asm volatile("bar.sync 1, 128;" ::: "memory");
            {
                int r = row;
                const int iw = r % g.bw; r /= g.bw;
                const int ih = r % g.bh; r /= g.bh;
                const int id = r % g.bd; r /= g.bd;
                row_valid[row] = (row < g.box_rows && w0 + iw < g.Wo && h0 + ih < g.Ho && d0 + id < g.Do && n0 + r < g.N) ? 1 : 0;
            }
            tc::mbar_wait(&tmem_full[buf], aph);
            tc::tc_fence_after();
            drain_tile<BLOCK_N>(tmem_base + buf * BLOCK_N + ((uint32_t)(q * 32) << 16), bias != nullptr ? bias_s : nullptr, staging, row);
            tc::tc_fence_before();
            tc::mbar_arrive(&tmem_empty[buf]);
            tc::fence_proxy_async_smem();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
#pragma unroll
                for (int j = 0; j < BLOCK_N / 64; ++j)
                    if (co0 + j * 64 < g.Cout) tc::tma_store_5d(&tm_y, staging + j * kATileBytes, co0 + j * 64, w0, h0, d0, n0);
                tc::tma_store_commit();
            }
            if (stats_partials != nullptr) {
                // per-channel (sum, sum^2) over the valid rows of this tile, from the bf16 values just staged
                float s1, s2;
                tile_stats<BLOCK_N>(staging, row_valid, g.box_rows, stat_red, et, s1, s2);
                if (et < BLOCK_N && co0 + et < g.Cout) {
                    const int m_tile = tile / g.n_tiles;
                    stats_partials[((size_t)m_tile * 2 + 0) * g.Cout + co0 + et] = s1;
                    stats_partials[((size_t)m_tile * 2 + 1) * g.Cout + co0 + et] = s2;
                }
            }
        }
        if (et == 0) tc::tma_store_wait_read();
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc<Cfg::kTmemCols>(tmem_base);
    }
}


// ------------------------------------------------------------------------------- fprop on CTA pairs (cta_group::2)
// Two CTAs of a cluster (one TPC) work on two neighbouring M tiles with ONE weight tile between them: every tcgen05.mma is
// M = 256 (128 positions from each CTA's shared memory) x N = 128, and each CTA stages only 64 of the 128 weight rows per
// K block -- the tensor cores of both SMs read the halves out of both shared memories.  Per CTA and K block that is 24 KiB
// instead of 32 KiB of L2 -> SM traffic and half the weight-operand reads per MMA (the operand-read bound of the 1-CTA
// N = 128 shape, scripts/bench_native/mma_rate.cu).  CTA 0 of the pair issues the MMAs; both run producers and epilogues.
// All TMA loads are plain loads of this GPU's activations / weights into the loading CTA's own shared memory (no multicast).
struct Fprop2Cfg {
    static constexpr int kBlockN = 128;
    static constexpr int kStages = 6;
    static constexpr int kBHalfBytes = 64 * 128;                 // 64 weight rows x 64 bf16
    static constexpr int kStageBytes = kATileBytes + kBHalfBytes;
    static constexpr int kStagingBytes = 2 * kATileBytes;
    static constexpr int kTmemCols = 256;
    static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 + 256 + 128 + 512 + kStatsRedBytes;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kConvThreads, 1) conv3d_fprop_tc2_kernel(
    const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_y,
    const ConvGeom g, const float* __restrict__ bias, float* __restrict__ stats_partials, int m_tiles, int total_items) {
    using Cfg = Fprop2Cfg;
    constexpr int BLOCK_N = Cfg::kBlockN;
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint8_t* stage_base = smem;
    uint8_t* staging = smem + Cfg::kStages * Cfg::kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
    uint64_t* full = bars;                       // [kStages]  (used in the pair's CTA 0)
    uint64_t* empty = bars + Cfg::kStages;       // [kStages]
    uint64_t* tmem_full = empty + Cfg::kStages;  // [2]
    uint64_t* tmem_empty = tmem_full + 2;        // [2]        (used in the pair's CTA 0: one arrival per CTA)
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    uint8_t* row_valid = reinterpret_cast<uint8_t*>(tmem_ptr + 2);   // [128]
    float* bias_s = reinterpret_cast<float*>(row_valid + 128);        // [128]
    float* stat_red = bias_s + 128;                                    // kStatsRedBytes

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const uint32_t rank = tc::cluster_ctarank();
    const int pair = (int)(blockIdx.x >> 1), n_pairs = (int)(gridDim.x >> 1);
    const int taps = g.k * g.k * g.k;
    const int num_kb = taps * g.cin_blocks;

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&tm_x);
        tc::tma_prefetch_desc(&tm_w);
        tc::tma_prefetch_desc(&tm_y);
        for (int i = 0; i < Cfg::kStages; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&tmem_full[i], 1); tc::mbar_init(&tmem_empty[i], 2); }
        tc::mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc_pair<Cfg::kTmemCols>(tmem_ptr);
    tc::tc_fence_before();
    tc::cluster_sync_all();          // barriers of both CTAs are initialised before anything arrives on them from the other CTA
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    // item = (pair of M tiles, N tile); this CTA's M tile is 2*pm + rank (clamped: an odd tile count leaves the last pair's
    // second CTA a duplicate that is computed but not stored)
    auto decode = [&](int item, int& n_tile, int& m_tile, bool& valid, int& w0, int& h0, int& d0, int& n0) {
        n_tile = item % g.n_tiles;
        int m = 2 * (item / g.n_tiles) + (int)rank;
        valid = m < m_tiles;
        if (!valid) m = m_tiles - 1;
        m_tile = m;
        w0 = (m % g.tw) * g.bw; m /= g.tw;
        h0 = (m % g.th) * g.bh; m /= g.th;
        d0 = (m % g.td) * g.bd; m /= g.td;
        n0 = m * g.bn;
    };

    if (warp == 0) {
        // ===== TMA producer (both CTAs): own A tile + own half of the weight tile; bytes are counted on CTA 0's full[s] =====
        const bool leader = tc::elect_one();
        uint32_t s = 0, ph = 0;
        for (int item = pair; item < total_items; item += n_pairs) {
            int n_tile, m_tile, w0, h0, d0, n0;
            bool valid;
            decode(item, n_tile, m_tile, valid, w0, h0, d0, n0);
            const int x0 = w0 * g.stride - g.pad, y0 = h0 * g.stride - g.pad, z0 = d0 * g.stride - g.pad;
            int kb = 0;
            for (int kd = 0; kd < g.k; ++kd)
                for (int kh = 0; kh < g.k; ++kh)
                    for (int kw = 0; kw < g.k; ++kw)
                        for (int cb = 0; cb < g.cin_blocks; ++cb, ++kb) {
                            tc::mbar_wait(&empty[s], ph ^ 1);
                            if (leader) {
                                uint8_t* a_dst = stage_base + s * Cfg::kStageBytes;
                                if (rank == 0) tc::mbar_expect_tx(&full[s], (uint32_t)(2 * (g.box_rows * 128 + Cfg::kBHalfBytes)));
                                tc::tma_load_5d_pair(&tm_x, &full[s], a_dst, cb * 64, x0 + kw, y0 + kh, z0 + kd, n0);
                                tc::tma_load_2d_pair(&tm_w, &full[s], a_dst + kATileBytes, kb * 64, n_tile * BLOCK_N + (int)rank * 64);
                            }
                            if (++s == (uint32_t)Cfg::kStages) { s = 0; ph ^= 1; }
                        }
        }
    } else if (warp == 1 && rank == 0) {
        // ===== MMA issuer (CTA 0 of the pair) =====
        const bool leader = tc::elect_one();
        constexpr uint32_t idesc = tc::make_idesc_bf16(256, BLOCK_N, 0, 0);
        constexpr uint32_t kDescHi = tc::desc_hi_sw128(1024);
        const uint32_t stage_lo = tc::desc_lo_sw128(tc::smem_u32(stage_base));
        uint32_t s = 0, ph = 0, local = 0;
        if (pair < total_items) {
            tc::mbar_wait(&full[0], 0);
            tc::tc_fence_after();
        }
        for (int item = pair; item < total_items; item += n_pairs, ++local) {
            const int buf = local & 1;
            tc::mbar_wait(&tmem_empty[buf], ((local >> 1) & 1) ^ 1);
            tc::tc_fence_after();
            const uint32_t d_tmem = tmem_base + buf * BLOCK_N;
            const bool more = item + n_pairs < total_items;
            for (int kb = 0; kb < num_kb; ++kb) {
                const uint32_t a_lo = stage_lo + s * (Cfg::kStageBytes >> 4);
                const uint32_t b_lo = a_lo + (kATileBytes >> 4);
                uint32_t s1 = s + 1, ph1 = ph;
                if (s1 == (uint32_t)Cfg::kStages) { s1 = 0; ph1 ^= 1; }
                if (leader) {
                    tc::umma_bf16_lohi_pair(d_tmem, a_lo, b_lo, kDescHi, idesc, kb != 0);
                    tc::umma_bf16_lohi_pair(d_tmem, a_lo + 2, b_lo + 2, kDescHi, idesc, 1);
                }
                if (kb + 1 < num_kb || more) {
                    tc::mbar_wait(&full[s1], ph1);
                    tc::tc_fence_after();
                }
                if (leader) {
                    tc::umma_bf16_lohi_pair(d_tmem, a_lo + 4, b_lo + 4, kDescHi, idesc, 1);
                    tc::umma_bf16_lohi_pair(d_tmem, a_lo + 6, b_lo + 6, kDescHi, idesc, 1);
                    tc::umma_commit_pair(&empty[s]);                 // frees stage s in both CTAs
                }
                s = s1; ph = ph1;
            }
            if (leader) tc::umma_commit_pair(&tmem_full[buf]);       // both epilogues
        }
    } else if (warp >= 2) {
        // ===== epilogue (both CTAs): this CTA's 128 rows of the pair's accumulator =====
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int et = threadIdx.x - 64;     // 0..127
        const uint32_t tmem_empty0 = tc::map_to_cta(tmem_empty, 0);        // CTA 0's barriers (shared::cluster address)
        uint32_t local = 0;
        for (int item = pair; item < total_items; item += n_pairs, ++local) {
            int n_tile, m_tile, w0, h0, d0, n0;
            bool valid;
            decode(item, n_tile, m_tile, valid, w0, h0, d0, n0);
            const int buf = local & 1;
            const uint32_t aph = (local >> 1) & 1;
            if (et == 0) tc::tma_store_wait_read();
            const int co0 = n_tile * BLOCK_N;
            if (bias != nullptr) bias_s[et] = (co0 + et < g.Cout) ? __ldg(bias + co0 + et) : 0.f;
            asm volatile("bar.sync 1, 128;" ::: "memory");
            {
                int r = row;
                const int iw = r % g.bw; r /= g.bw;
                const int ih = r % g.bh; r /= g.bh;
                const int id = r % g.bd; r /= g.bd;
                row_valid[row] = (row < g.box_rows && w0 + iw < g.Wo && h0 + ih < g.Ho && d0 + id < g.Do && n0 + r < g.N) ? 1 : 0;
            }
            tc::mbar_wait(&tmem_full[buf], aph);
            tc::tc_fence_after();
            drain_tile<BLOCK_N>(tmem_base + buf * BLOCK_N + ((uint32_t)(q * 32) << 16), bias != nullptr ? bias_s : nullptr, staging, row);
            tc::tc_fence_before();
            tc::fence_proxy_async_smem();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (et == 0) {
                tc::mbar_arrive_cluster(tmem_empty0 + (uint32_t)(buf * sizeof(uint64_t)));   // this CTA has drained its half of buffer `buf`
                if (valid) {
#pragma unroll
                    for (int j = 0; j < BLOCK_N / 64; ++j)
                        if (co0 + j * 64 < g.Cout) tc::tma_store_5d(&tm_y, staging + j * kATileBytes, co0 + j * 64, w0, h0, d0, n0);
                }
                tc::tma_store_commit();
            }
            if (stats_partials != nullptr) {        // (uniform over the CTA: the barrier inside is reached by all 128 threads)
                float s1, s2;
                tile_stats<BLOCK_N>(staging, row_valid, g.box_rows, stat_red, et, s1, s2);
                if (valid && co0 + et < g.Cout) {
                    stats_partials[((size_t)m_tile * 2 + 0) * g.Cout + co0 + et] = s1;
                    stats_partials[((size_t)m_tile * 2 + 1) * g.Cout + co0 + et] = s2;
                }
            }
        }
        if (et == 0) tc::tma_store_wait_read();
    }
    tc::tc_fence_before();
    tc::cluster_sync_all();          // neither CTA leaves (or frees TMEM) while the other may still touch its barriers / operands
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc_pair<Cfg::kTmemCols>(tmem_base);
    }
}

// ------------------------------------------------------------------------------- wgrad
// Work item = (tap, co tile of 128, ci block of 64, K split).  D[co][ci] (128x64 fp32 in TMEM)
// accumulates dY^T * X_shift over the positions of the split; both operands are MN-major
// (positions are the GEMM K dimension, channels are contiguous in memory).
struct WgradGeom {
    int N, Do, Ho, Wo, Cout, Cin;
    int k, stride, pad;
    int bw, bh, bd, bn, tw, th, td, tn;
    int box_rows, m_tiles, splits, co_tiles, ci_blocks;
};
constexpr int kWgStages = 4;
constexpr int kWgStageBytes = 3 * kATileBytes;    // A: 2 x [128 pos][64 co], B: [128 pos][64 ci]
constexpr int kWgSmemBytes = kWgStages * kWgStageBytes + 1024 + 256;

__global__ void __launch_bounds__(kConvThreads, 1) conv3d_wgrad_tc_kernel(
    const __grid_constant__ CUtensorMap tm_dy, const __grid_constant__ CUtensorMap tm_x, const WgradGeom g, float* __restrict__ partials) {
    extern __shared__ __align__(1024) uint8_t smem[];   // SWIZZLE_128B tiles need 1024-byte alignment
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kWgStages * kWgStageBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + kWgStages;
    uint64_t* tmem_full = empty + kWgStages;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;     // warp-uniform (see fprop)
    const int taps = g.k * g.k * g.k;

    // work item decode: ci block fastest, then co tile, then tap, then split
    int wi = blockIdx.x;
    const int cib = wi % g.ci_blocks; wi /= g.ci_blocks;
    const int cot = wi % g.co_tiles; wi /= g.co_tiles;
    const int tap = wi % taps;
    const int split = wi / taps;
    const int per = (g.m_tiles + g.splits - 1) / g.splits;
    const int mt_begin = split * per;
    const int mt_end = min(g.m_tiles, mt_begin + per);
    const int kd = tap / (g.k * g.k), kh = (tap / g.k) % g.k, kw = tap % g.k;

    // rows >= box_rows of every stage are never written by TMA: keep them zero (they are GEMM-K here)
    for (int i = threadIdx.x; i < kWgStages * kWgStageBytes / 16; i += kConvThreads) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    tc::fence_proxy_async_smem();
    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&tm_dy);
        tc::tma_prefetch_desc(&tm_x);
        for (int i = 0; i < kWgStages; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
        tc::mbar_init(tmem_full, 1);
        tc::mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc<64>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        const bool leader = tc::elect_one();
        uint32_t s = 0, ph = 0;
        // running tile coordinates instead of a division chain per tile
        int m = mt_begin;
        int iw = m % g.tw; m /= g.tw;
        int ih = m % g.th; m /= g.th;
        int id = m % g.td;
        int in = m / g.td;
        for (int mt = mt_begin; mt < mt_end; ++mt) {
            const int w0 = iw * g.bw, h0 = ih * g.bh, d0 = id * g.bd, n0 = in * g.bn;
            tc::mbar_wait(&empty[s], ph ^ 1);
            if (leader) {
                uint8_t* a_dst = smem + s * kWgStageBytes;
                tc::mbar_expect_tx(&full[s], (uint32_t)(3 * g.box_rows * 128));
                tc::tma_load_5d(&tm_dy, &full[s], a_dst, cot * 128, w0, h0, d0, n0);
                tc::tma_load_5d(&tm_dy, &full[s], a_dst + kATileBytes, cot * 128 + 64, w0, h0, d0, n0);
                tc::tma_load_5d(&tm_x, &full[s], a_dst + 2 * kATileBytes, cib * 64, w0 * g.stride - g.pad + kw, h0 * g.stride - g.pad + kh,
                                d0 * g.stride - g.pad + kd, n0);
            }
            if (++s == (uint32_t)kWgStages) { s = 0; ph ^= 1; }
            if (++iw == g.tw) { iw = 0; if (++ih == g.th) { ih = 0; if (++id == g.td) { id = 0; ++in; } } }
        }
    } else if (warp == 1) {
        const bool leader = tc::elect_one();
        constexpr uint32_t idesc = tc::make_idesc_bf16(128, 64, 1, 1);
        constexpr uint32_t kDescHi = tc::desc_hi_sw128(1024);
        const uint32_t stage_lo = tc::desc_lo_sw128(tc::smem_u32(smem), kATileBytes);     // MN-major: LBO = the 64-channel half (16 KB)
        uint32_t s = 0, ph = 0;
        if (mt_begin < mt_end) {
            tc::mbar_wait(&full[0], 0);
            tc::tc_fence_after();
        }
        for (int mt = mt_begin; mt < mt_end; ++mt) {
            const uint32_t a_lo = stage_lo + s * (kWgStageBytes >> 4);
            const uint32_t b_lo = a_lo + (2 * kATileBytes >> 4);
            uint32_t s1 = s + 1, ph1 = ph;
            if (s1 == (uint32_t)kWgStages) { s1 = 0; ph1 ^= 1; }
            if (leader) {
#pragma unroll
                for (int k = 0; k < 4; ++k)     // 16 positions per MMA: two 8-row swizzle groups (SBO = 1024 B); +2048 B per step
                    tc::umma_bf16_lohi(tmem_base, a_lo + k * 128, b_lo + k * 128, kDescHi, idesc, (k != 0) || (mt != mt_begin));
            }
            if (mt + 1 < mt_end) {              // next stage's wait in the middle of this stage's MMAs (see fprop)
                tc::mbar_wait(&full[s1], ph1);
                tc::tc_fence_after();
            }
            if (leader) {
#pragma unroll
                for (int k = 4; k < 8; ++k)
                    tc::umma_bf16_lohi(tmem_base, a_lo + k * 128, b_lo + k * 128, kDescHi, idesc, 1);
                tc::umma_commit(&empty[s]);
            }
            s = s1; ph = ph1;
        }
        if (leader) tc::umma_commit(tmem_full);
    } else {
        const int q = warp & 3;
        const int co = cot * 128 + q * 32 + lane;
        if (mt_end > mt_begin) {
            tc::mbar_wait(tmem_full, 0);
            tc::tc_fence_after();
        }
        float* dst = partials + (((size_t)split * taps + tap) * g.Cout + co) * g.Cin + cib * 64;
#pragma unroll 1
        for (int c = 0; c < 64; c += 16) {
            float v[16];
            if (mt_end > mt_begin) tc::tmem_ld16(tmem_base + c + ((uint32_t)(q * 32) << 16), v);
            else {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = 0.f;
            }
            if (co < g.Cout) {
#pragma unroll
                for (int i = 0; i < 16; i += 4) st4(dst + c + i, make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
            }
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc<64>(tmem_base);
    }
}

// dW (PyTorch layout [co][ci][taps], fp32, in the gradient slab) = sum_splits partials[split][tap][co][ci];
// optional saliency epilogue: scores[co][ci][tap] += scale * |dW * W|   (ref K11: snip.py:60-71)
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partials, int splits, int taps, int Cout, int Cin,
                                                           float* __restrict__ grad, const float* __restrict__ w, float* __restrict__ scores,
                                                           float scale) {
    // threads follow the partial layout [tap][co][ci] (ci fastest): coalesced reads of every split, strided writes of the result
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cin);
        const int64_t r = i / Cin;
        const int co = (int)(r % Cout);
        const int t = (int)(r / Cout);
        float acc = 0.f;
        for (int s = 0; s < splits; ++s) acc += partials[(size_t)s * total + i];
        const size_t o = ((size_t)co * Cin + ci) * taps + t;
        grad[o] = acc;
        if (scores != nullptr) scores[o] = fmaf(scale, fabsf(acc * w[o]), scores[o]);      // same arithmetic as saliency_accumulate_kernel
    }
}

// fp32 master weights [co][ci][taps] -> bf16 GEMM operands:
//   wf[co][tap][ci]            (fprop B operand, K-major)
//   wd[ci][taps-1-tap][co]     (dgrad B operand: transposed and tap-flipped)
__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ w, int Cout, int Cin, int taps,
                                                           __nv_bfloat16* __restrict__ wf, __nv_bfloat16* __restrict__ wd) {
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        // iterate in wf order so the writes of the larger-stride operand stay coalesced
        const int ci = (int)(i % Cin);
        const int64_t r = i / Cin;
        const int t = (int)(r % taps);
        const int co = (int)(r / taps);
        const __nv_bfloat16 v = __float2bfloat16(w[((size_t)co * Cin + ci) * taps + t]);
        wf[i] = v;
        if (wd != nullptr) wd[((size_t)ci * taps + (taps - 1 - t)) * Cout + co] = v;
    }
}

// all conv layers of a model in ONE launch (a step re-packs every layer after the SGD update; four ~10 us launches otherwise)
struct PackMulti {
    const float* w[8];
    __nv_bfloat16* wf[8];
    __nv_bfloat16* wd[8];
    int Cout[8], Cin[8], taps[8];
    int block_end[8];       // exclusive prefix of blocks per layer
    int n;
};
__global__ void __launch_bounds__(256) pack_weights_multi_kernel(const PackMulti p) {
    int l = 0;
    while (l + 1 < p.n && (int)blockIdx.x >= p.block_end[l]) ++l;
    const int b0 = l == 0 ? 0 : p.block_end[l - 1];
    const int nb = p.block_end[l] - b0;
    const int Cout = p.Cout[l], Cin = p.Cin[l], taps = p.taps[l];
    const float* __restrict__ w = p.w[l];
    __nv_bfloat16* __restrict__ wf = p.wf[l];
    __nv_bfloat16* __restrict__ wd = p.wd[l];
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)(blockIdx.x - b0) * blockDim.x + threadIdx.x; i < total; i += (int64_t)nb * blockDim.x) {
        const int ci = (int)(i % Cin);
        const int64_t r = i / Cin;
        const int t = (int)(r % taps);
        const int co = (int)(r / taps);
        const __nv_bfloat16 v = __float2bfloat16(w[((size_t)co * Cin + ci) * taps + t]);
        wf[i] = v;
        if (wd != nullptr) wd[((size_t)ci * taps + (taps - 1 - t)) * Cout + co] = v;
    }
}

}  // namespace nidt
using namespace nidt;

// n <= 8 layers; arrays are HOST arrays of device pointers / shapes
NIDT_API int nidt_pack_conv_weights_multi(int n, const float* const* w, const int* Cout, const int* Cin, const int* taps, void* const* wf,
                                          void* const* wd, void* stream) {
    NIDT_REQUIRE(n >= 1 && n <= 8, "nidt_pack_conv_weights_multi: %d layers (1..8)", n);
    PackMulti p;
    int blocks = 0;
    for (int l = 0; l < n; ++l) {
        p.w[l] = w[l]; p.wf[l] = (__nv_bfloat16*)wf[l]; p.wd[l] = (__nv_bfloat16*)wd[l];
        p.Cout[l] = Cout[l]; p.Cin[l] = Cin[l]; p.taps[l] = taps[l];
        int64_t nb = ((int64_t)Cout[l] * Cin[l] * taps[l] + 255) / 256;
        if (nb > kNumSMs * 4) nb = kNumSMs * 4;
        blocks += (int)nb;
        p.block_end[l] = blocks;
    }
    p.n = n;
    pack_weights_multi_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(p);
    NIDT_LAUNCH_CHECK();
    return 0;
}

static int pick_tiles(int n, int b) { return (n + b - 1) / b; }

// x: bf16 [N][Di][Hi][Wi][Cin]; w: bf16 [Cout][k^3][Cin]; y: bf16 [N][Do][Ho][Wo][Cout]
// stats_partials: nullptr or fp32 [m_tiles][2][Cout]; box (bw,bh,bd,bn) with bw*bh*bd*bn <= 128.
NIDT_API int nidt_conv3d_fprop_tc(const void* x, const void* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                                  int Cin, int Do, int Ho, int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn,
                                  int block_n, void* stream) {
    NIDT_REQUIRE(Cin % 64 == 0 && Cout % 8 == 0, "conv3d_fprop_tc: Cin=%d must be a multiple of 64, Cout=%d of 8", Cin, Cout);
    NIDT_REQUIRE(bw * bh * bd * bn <= 128 && bw >= 1 && bh >= 1 && bd >= 1 && bn >= 1, "conv3d_fprop_tc: box %dx%dx%dx%d exceeds 128 rows", bw, bh, bd, bn);
    NIDT_REQUIRE(block_n == 64 || block_n == 128, "conv3d_fprop_tc: block_n must be 64 or 128");
    ConvGeom g;
    g.N = N; g.Do = Do; g.Ho = Ho; g.Wo = Wo; g.Cout = Cout; g.k = k; g.stride = stride; g.pad = pad; g.cin_blocks = Cin / 64;
    g.bw = bw; g.bh = bh; g.bd = bd; g.bn = bn;
    g.tw = pick_tiles(Wo, bw); g.th = pick_tiles(Ho, bh); g.td = pick_tiles(Do, bd); g.tn = pick_tiles(N, bn);
    g.n_tiles = pick_tiles(Cout, block_n);
    g.box_rows = bw * bh * bd * bn;
    const int m_tiles = g.tw * g.th * g.td * g.tn;
    const int total = m_tiles * g.n_tiles;
    CUtensorMap tx, tw_, ty;
    if (make_tmap_act5d(&tx, x, N, Di, Hi, Wi, Cin, bw, bh, bd, bn, stride)) return 1;
    if (make_tmap_mat2d(&tw_, w, Cout, (int64_t)k * k * k * Cin, block_n)) return 1;
    if (make_tmap_act5d(&ty, y, N, Do, Ho, Wo, Cout, bw, bh, bd, bn, 1)) return 1;
    int dev = 0, sms = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int grid = total < sms ? total : sms;
    cudaStream_t s = (cudaStream_t)stream;
    static const int use_pairs = [] { const char* e = getenv("NIDT_CONV_PAIRS"); return e ? atoi(e) : 1; }();
    if (block_n == 128 && use_pairs && m_tiles >= 2) {
        // CTA pairs: two M tiles share one weight tile through tcgen05.mma.cta_group::2 (cluster of 2 CTAs = one TPC)
        CUtensorMap tw_half;
        if (make_tmap_mat2d(&tw_half, w, Cout, (int64_t)k * k * k * Cin, 64)) return 1;
        const int items = ((m_tiles + 1) / 2) * g.n_tiles;
        int pairs = sms / 2;
        if (pairs > items) pairs = items;
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv3d_fprop_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Fprop2Cfg::kSmemBytes));
        conv3d_fprop_tc2_kernel<<<2 * pairs, kConvThreads, Fprop2Cfg::kSmemBytes, s>>>(tx, tw_half, ty, g, bias, stats_partials, m_tiles, items);
        NIDT_LAUNCH_CHECK();
        return 0;
    }
    if (block_n == 128) {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv3d_fprop_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, FpropCfg<128>::kSmemBytes));
        conv3d_fprop_tc_kernel<128><<<grid, kConvThreads, FpropCfg<128>::kSmemBytes, s>>>(tx, tw_, ty, g, bias, stats_partials, total);
    } else {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv3d_fprop_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, FpropCfg<64>::kSmemBytes));
        conv3d_fprop_tc_kernel<64><<<grid, kConvThreads, FpropCfg<64>::kSmemBytes, s>>>(tx, tw_, ty, g, bias, stats_partials, total);
    }
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_conv3d_m_tiles(int N, int Do, int Ho, int Wo, int bw, int bh, int bd, int bn) {
    return pick_tiles(Wo, bw) * pick_tiles(Ho, bh) * pick_tiles(Do, bd) * pick_tiles(N, bn);
}

NIDT_API int64_t nidt_conv3d_wgrad_scratch_floats(int Cout, int Cin, int k, int splits) { return (int64_t)splits * k * k * k * Cout * Cin; }

static int wgrad_impl(const void* dy, const void* x, float* grad, float* scratch, int N, int Di, int Hi, int Wi, int Cin, int Do, int Ho,
                      int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn, int splits, void* stream,
                      const float* sal_w = nullptr, float* sal_scores = nullptr, float sal_scale = 0.f) {
    NIDT_REQUIRE(Cin % 64 == 0 && Cout % 8 == 0, "conv3d_wgrad_tc: Cin=%d must be a multiple of 64, Cout=%d of 8", Cin, Cout);
    NIDT_REQUIRE(bw * bh * bd * bn <= 128, "conv3d_wgrad_tc: box exceeds 128 rows");
    NIDT_REQUIRE(splits >= 1 && scratch != nullptr, "conv3d_wgrad_tc: need splits >= 1 and a scratch buffer");
    WgradGeom g;
    g.N = N; g.Do = Do; g.Ho = Ho; g.Wo = Wo; g.Cout = Cout; g.Cin = Cin; g.k = k; g.stride = stride; g.pad = pad;
    g.bw = bw; g.bh = bh; g.bd = bd; g.bn = bn;
    g.tw = pick_tiles(Wo, bw); g.th = pick_tiles(Ho, bh); g.td = pick_tiles(Do, bd); g.tn = pick_tiles(N, bn);
    g.box_rows = bw * bh * bd * bn;
    g.m_tiles = g.tw * g.th * g.td * g.tn;
    g.splits = splits;
    g.co_tiles = pick_tiles(Cout, 128);
    g.ci_blocks = Cin / 64;
    const int taps = k * k * k;
    CUtensorMap tdy, tx;
    if (make_tmap_act5d(&tdy, dy, N, Do, Ho, Wo, Cout, bw, bh, bd, bn, 1)) return 1;
    if (make_tmap_act5d(&tx, x, N, Di, Hi, Wi, Cin, bw, bh, bd, bn, stride)) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv3d_wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmemBytes));
    const int grid = splits * taps * g.co_tiles * g.ci_blocks;
    conv3d_wgrad_tc_kernel<<<grid, kConvThreads, kWgSmemBytes, s>>>(tdy, tx, g, scratch);
    NIDT_LAUNCH_CHECK();
    const int64_t total = (int64_t)Cout * Cin * taps;
    int64_t blocks = (total + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    wgrad_reduce_kernel<<<(int)blocks, 256, 0, s>>>(scratch, splits, taps, Cout, Cin, grad, sal_w, sal_scores, sal_scale);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// dy: bf16 [N][Do][Ho][Wo][Cout]; x: bf16 [N][Di][Hi][Wi][Cin]; grad: fp32 [Cout][Cin][k^3] (PyTorch layout);
// scratch: >= nidt_conv3d_wgrad_scratch_floats() floats of split-K partials.
NIDT_API int nidt_conv3d_wgrad_tc(const void* dy, const void* x, float* grad, float* scratch, int N, int Di, int Hi, int Wi, int Cin, int Do, int Ho,
                                  int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn, int splits, void* stream) {
    return wgrad_impl(dy, x, grad, scratch, N, Di, Hi, Wi, Cin, Do, Ho, Wo, Cout, k, stride, pad, bw, bh, bd, bn, splits, stream);
}

// The same weight gradient with the SNIP saliency epilogue: the split-K reduction that writes dW also does
// scores[co][ci][tap] += scale * |dW * w| (w: the layer's fp32 master weights, same layout) -- the score leaves the wgrad
// reduction directly instead of a later pass over the gradient slab.  ref: snip.py:60-71 (|weight_mask.grad| = |dL/dw * w|).
NIDT_API int nidt_conv3d_wgrad_tc_saliency(const void* dy, const void* x, float* grad, float* scratch, int N, int Di, int Hi, int Wi, int Cin, int Do,
                                           int Ho, int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn, int splits,
                                           const float* w, float* scores, float scale, void* stream) {
    NIDT_REQUIRE(w != nullptr && scores != nullptr, "conv3d_wgrad_tc_saliency: w / scores is null");
    return wgrad_impl(dy, x, grad, scratch, N, Di, Hi, Wi, Cin, Do, Ho, Wo, Cout, k, stride, pad, bw, bh, bd, bn, splits, stream, w, scores, scale);
}

NIDT_API int nidt_pack_conv_weights(const float* w, int Cout, int Cin, int taps, void* wf, void* wd, void* stream) {
    const int64_t total = (int64_t)Cout * Cin * taps;
    int64_t blocks = (total + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    pack_weights_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(w, Cout, Cin, taps, (__nv_bfloat16*)wf, (__nv_bfloat16*)wd);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------------
// nidt_conv3d_wgrad_tc_static: the same weight-gradient kernels, but the split-K partials live in a
// scratch array that is part of the library image instead of a buffer the caller allocates.
//
// It is meant for callers that hold device pointers but have no way to obtain more device memory at
// the point of the call: a custom-call body inside another framework's executor (which owns the
// allocator and hands the callee only its operands), and deployments whose memory is carved up once at
// start-up and where a late cudaMalloc is forbidden or may fail.  Such a caller passes only the inputs
// (dy, x) and the output (grad).
//
// Cost and limits: kWgradStaticFloats floats (16 MiB) of global memory are held on every GPU on which the
// library's device code gets loaded (with CUDA's lazy module loading: from the first use of the library
// on that GPU, at the latest the cudaGetSymbolAddress below) until the process exits.
// A problem needs splits*k^3*Cout*Cin floats; anything larger is refused (return code 2, message via
// nidt_last_error()) -- use nidt_conv3d_wgrad_tc with your own buffer for those.
//
// Set-up: the first call on each device resolves the device address of the array with
// cudaGetSymbolAddress (the symbol's name is only a host-side handle, not something a kernel can be
// given) and caches it.  Make that first call during initialisation, with the target device current,
// before starting any stream capture.  No cudaMalloc / cudaFree / allocating library routine is called.
//
// Sharing: all calls on a device use the same array and it keeps whatever the previous call left there.
// Every partial that a call reads back is first written by that call's own GEMM kernel (empty splits
// write zeros), so nothing needs clearing between calls -- but two calls whose kernels may overlap on the
// GPU (different streams, or graph branches without a dependency) would overwrite each other's partials:
// serialise them.
constexpr size_t kWgradStaticFloats = (size_t)4 << 20;
static __device__ __align__(256) float g_wgrad_static_scratch[kWgradStaticFloats];

constexpr int kMaxStaticDevices = 64;
static float* g_wgrad_static_ptr[kMaxStaticDevices];   // device address of the array, per device ordinal
static std::mutex g_wgrad_static_mu;

NIDT_API int64_t nidt_conv3d_wgrad_static_capacity_floats() { return (int64_t)kWgradStaticFloats; }

NIDT_API int nidt_conv3d_wgrad_tc_static(const void* dy, const void* x, float* grad, int N, int Di, int Hi, int Wi, int Cin, int Do, int Ho,
                                         int Wo, int Cout, int k, int stride, int pad, int bw, int bh, int bd, int bn, int splits, void* stream) {
    NIDT_REQUIRE(splits >= 1 && k >= 1 && Cout >= 1 && Cin >= 1, "conv3d_wgrad_tc_static: bad shape (splits=%d k=%d Cout=%d Cin=%d)", splits, k, Cout, Cin);
    const int64_t need = (int64_t)splits * k * k * k * Cout * Cin;
    NIDT_REQUIRE(need <= (int64_t)kWgradStaticFloats,
                 "conv3d_wgrad_tc_static: needs %lld scratch floats (splits=%d x k^3=%d x Cout=%d x Cin=%d) but the built-in scratch holds %lld; "
                 "lower `splits` or call nidt_conv3d_wgrad_tc with a caller-provided buffer",
                 (long long)need, splits, k * k * k, Cout, Cin, (long long)kWgradStaticFloats);
    int dev = -1;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_REQUIRE(dev >= 0 && dev < kMaxStaticDevices, "conv3d_wgrad_tc_static: device ordinal %d out of range", dev);
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_wgrad_static_mu);
        if (g_wgrad_static_ptr[dev] == nullptr) {
            void* p = nullptr;
            NIDT_CUDA_CHECK(cudaGetSymbolAddress(&p, g_wgrad_static_scratch));
            g_wgrad_static_ptr[dev] = static_cast<float*>(p);
        }
        scratch = g_wgrad_static_ptr[dev];
    }
    return wgrad_impl(dy, x, grad, scratch, N, Di, Hi, Wi, Cin, Do, Ho, Wo, Cout, k, stride, pad, bw, bh, bd, bn, splits, stream);
}
