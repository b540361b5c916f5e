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
// fprop's epilogue adds the bias, stores bf16 NDHWC tiles by TMA and accumulates the BatchNorm
// (sum, sum^2) partials per persistent CTA.  Reference op: salient_models.py:147 (K1/K2, cuDNN fp32).
#include "tc_common.cuh"

namespace nidt {

struct InConvGeom {
    int N, Di, Hi, Wi;      // input volume (1 channel, u8)
    int Do, Ho, Wo;         // output extents
    int stride, pad;
    int TH;                 // output rows (h) per tile: TH*Wo <= 128
    int h_tiles;            // ceil(Ho / TH)
    int PH, PW;             // patch rows = (TH-1)*stride + k ; row pitch in elements (even)
    int total_tiles;        // N * Do * h_tiles
    float scale;            // x = u8 * scale
};

constexpr int kInThreads = 13 * 32;   // warps 0-3 / 4-7: builder groups, warp 8: MMA, warps 9-12: epilogue (+TMA)
constexpr int kT16 = 16384;           // one [128 rows][64 bf16] swizzled sub-tile

__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// ---- stage the u8 input patch of one tile as normalised bf16: patch[KS][PH][PW]
template <int KS>
__device__ __forceinline__ void stage_patch(const uint8_t* __restrict__ x, __nv_bfloat16* patch, const InConvGeom& g, int n, int d, int h0, int gt) {
    const int total = KS * g.PH * g.PW;
    const int z0 = d * g.stride - g.pad, y0 = h0 * g.stride - g.pad;
    for (int idx = gt; idx < total; idx += 128) {
        const int w = idx % g.PW;
        const int t = idx / g.PW;
        const int hr = t % g.PH, kd = t / g.PH;
        const int iz = z0 + kd, iy = y0 + hr, ix = w - g.pad;
        float v = 0.f;
        if (iz >= 0 && iz < g.Di && iy >= 0 && iy < g.Hi && ix >= 0 && ix < g.Wi)
            v = (float)x[(((size_t)n * g.Di + iz) * g.Hi + iy) * g.Wi + ix] * g.scale;
        patch[idx] = __float2bfloat16(v);
    }
}

// ---- one im2col row (= one output position, 128 K slots) into the two swizzled sub-tiles of an A tile
template <int KS>
__device__ __forceinline__ void build_row(const __nv_bfloat16* patch, const InConvGeom& g, int valid_rows, uint8_t* a_tile, int r) {
    constexpr int kTaps = KS * KS * KS;
    const int sw = r & 7;
    uint8_t* row0 = a_tile + r * 128;
    if (r < valid_rows) {
        const int hh = r / g.Wo, w = r - hh * g.Wo;
        const unsigned short* base = reinterpret_cast<const unsigned short*>(patch) + (hh * g.stride) * g.PW + w * g.stride;
#pragma unroll
        for (int chunk = 0; chunk < 16; ++chunk) {
            uint32_t wv[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t lo = 0, hi = 0;
                const int e0 = chunk * 8 + 2 * j, e1 = e0 + 1;
                if (e0 < kTaps) lo = base[((e0 / (KS * KS)) * g.PH + (e0 / KS) % KS) * g.PW + e0 % KS];
                if (e1 < kTaps) hi = base[((e1 / (KS * KS)) * g.PH + (e1 / KS) % KS) * g.PW + e1 % KS];
                wv[j] = lo | (hi << 16);
            }
            *reinterpret_cast<uint4*>(row0 + (chunk >> 3) * kT16 + (((chunk & 7) ^ sw) << 4)) = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        }
    } else {
#pragma unroll
        for (int chunk = 0; chunk < 16; ++chunk)
            *reinterpret_cast<uint4*>(row0 + (chunk >> 3) * kT16 + (((chunk & 7) ^ sw) << 4)) = make_uint4(0, 0, 0, 0);
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

// =========================================================================== fprop
template <int KS>
__global__ void __launch_bounds__(kInThreads, 1) conv_in_fprop_kernel(
    const uint8_t* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, const __grid_constant__ CUtensorMap tm_y,
    float* __restrict__ stats_partials, const InConvGeom g) {
    constexpr int kTaps = KS * KS * KS;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* a_tiles = smem;                       // [2][2*kT16]
    uint8_t* b_tile = smem + 4 * kT16;             // [2 sub][64 rows][128 B]
    uint8_t* staging = b_tile + kT16;              // [128][128 B] swizzled
    uint64_t* bars = reinterpret_cast<uint64_t*>(staging + kT16);
    uint64_t *a_full = bars, *a_empty = bars + 2, *t_full = bars + 4, *t_empty = bars + 6;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 8);
    float* red = reinterpret_cast<float*>(tmem_ptr + 4);     // [2][2][64]
    __nv_bfloat16* patches = reinterpret_cast<__nv_bfloat16*>(red + 256);
    const int patch_elems = (KS * g.PH * g.PW + 7) & ~7;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // weights -> bf16 K-major swizzled B tile (rows = co)
    for (int idx = threadIdx.x; idx < 64 * 128; idx += kInThreads) {
        const int co = idx >> 7, kk = idx & 127;
        const float v = kk < kTaps ? w[co * kTaps + kk] : 0.f;
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
        // ===== builders =====
        const int grp = warp >> 2, gt = threadIdx.x & 127;
        __nv_bfloat16* patch = patches + grp * patch_elems;
        uint8_t* a_tile = a_tiles + grp * 2 * kT16;
        for (int i = grp;; i += 2) {
            const int tile = blockIdx.x + i * gridDim.x;
            if (tile >= g.total_tiles) break;
            const uint32_t u = i >> 1;
            int n, d, h0, valid;
            decode_tile(g, tile, n, d, h0, valid);
            tc::mbar_wait(&a_empty[grp], (u & 1) ^ 1);
            stage_patch<KS>(x, patch, g, n, d, h0, gt);
            named_bar(1 + grp, 128);
            build_row<KS>(patch, g, valid, a_tile, gt);
            tc::fence_proxy_async_smem();
            tc::mbar_arrive(&a_full[grp]);
        }
    } else if (warp == 8) {
        if (lane == 0) {
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
                    tc::umma_bf16(tmem_base + b * 64, da, db, idesc, k != 0);
                }
                tc::umma_commit(&a_empty[b]);
                tc::umma_commit(&t_full[b]);
            }
        }
    } else {
        // ===== epilogue =====
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int et = threadIdx.x - 9 * 32;
        const int ch = et & 63, half = et >> 6;
        float s1 = 0.f, s2 = 0.f;
        for (int i = 0;; ++i) {
            const int tile = blockIdx.x + i * gridDim.x;
            if (tile >= g.total_tiles) break;
            const int b = i & 1;
            const uint32_t u = i >> 1;
            int n, d, h0, valid;
            decode_tile(g, tile, n, d, h0, valid);
            if (et == 0) tc::tma_store_wait_read();
            named_bar(3, 128);
            tc::mbar_wait(&t_full[b], u & 1);
            tc::tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 64; c += 16) {
                float v[16];
                tc::tmem_ld16(tmem_base + b * 64 + c + ((uint32_t)(q * 32) << 16), v);
                if (bias != nullptr) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] += __ldg(bias + c + j);
                }
                uint4 o[2];
                __nv_bfloat162* ob = reinterpret_cast<__nv_bfloat162*>(o);
#pragma unroll
                for (int j = 0; j < 8; ++j) ob[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                uint8_t* blk = staging + row * 128;
                const int c0 = c >> 3;
                *reinterpret_cast<uint4*>(blk + ((c0 ^ (row & 7)) << 4)) = o[0];
                *reinterpret_cast<uint4*>(blk + (((c0 + 1) ^ (row & 7)) << 4)) = o[1];
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&t_empty[b]);
            tc::fence_proxy_async_smem();
            named_bar(3, 128);
            if (et == 0) {
                tc::tma_store_5d(&tm_y, staging, 0, 0, h0, d, n);
                tc::tma_store_commit();
            }
            if (stats_partials != nullptr) {
                const int chunk = ch >> 3, within = (ch & 7) * 2;
                for (int r = half; r < valid; r += 2) {
                    const float xv = __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(staging + r * 128 + ((chunk ^ (r & 7)) << 4) + within));
                    s1 += xv;
                    s2 = fmaf(xv, xv, s2);
                }
            }
        }
        if (et == 0) tc::tma_store_wait_read();
        if (stats_partials != nullptr) {
            red[(half * 2 + 0) * 64 + ch] = s1;
            red[(half * 2 + 1) * 64 + ch] = s2;
            named_bar(3, 128);
            if (half == 0) {
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
template <int KS>
__global__ void __launch_bounds__(kInThreads, 1) conv_in_wgrad_kernel(
    const uint8_t* __restrict__ x, const __grid_constant__ CUtensorMap tm_dy, float* __restrict__ partials /* [grid][128][64] */, const InConvGeom g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* a_tiles = smem;                   // [2][2*kT16]
    uint8_t* dy_tiles = smem + 4 * kT16;       // [2][kT16]
    uint64_t* bars = reinterpret_cast<uint64_t*>(dy_tiles + 2 * kT16);
    uint64_t *a_full = bars, *a_empty = bars + 2, *dy_full = bars + 4, *t_full = bars + 6;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 8);
    __nv_bfloat16* patches = reinterpret_cast<__nv_bfloat16*>(tmem_ptr + 4);
    const int patch_elems = (KS * g.PH * g.PW + 7) & ~7;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

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
        const int grp = warp >> 2, gt = threadIdx.x & 127;
        __nv_bfloat16* patch = patches + grp * patch_elems;
        uint8_t* a_tile = a_tiles + grp * 2 * kT16;
        for (int i = grp;; i += 2) {
            const int tile = blockIdx.x + i * gridDim.x;
            if (tile >= g.total_tiles) break;
            const uint32_t u = i >> 1;
            int n, d, h0, valid;
            decode_tile(g, tile, n, d, h0, valid);
            tc::mbar_wait(&a_empty[grp], (u & 1) ^ 1);
            stage_patch<KS>(x, patch, g, n, d, h0, gt);
            named_bar(1 + grp, 128);
            build_row<KS>(patch, g, valid, a_tile, gt);
            tc::fence_proxy_async_smem();
            tc::mbar_arrive(&a_full[grp]);
        }
    } else if (warp == 8) {
        if (lane == 0) {
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
                    tc::umma_bf16(tmem_base, da, db, idesc, (i | k) != 0);
                }
                tc::umma_commit(&a_empty[b]);
            }
            tc::umma_commit(t_full);
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

// dW[co][tap] = sum_cta partials[cta][tap][co]
__global__ void __launch_bounds__(256) conv_in_wgrad_reduce_kernel(const float* __restrict__ partials, int n_cta, int taps, float* __restrict__ grad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 64 * taps) return;
    const int co = i / taps, t = i % taps;
    float acc = 0.f;
    for (int c = 0; c < n_cta; ++c) acc += partials[((size_t)c * 128 + t) * 64 + co];
    grad[i] = acc;
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
    g.PW = (Wi + 2 * pad + 1) & ~1;
    g.total_tiles = N * g.Do * g.h_tiles;
    return 0;
}
static size_t in_smem_bytes(const InConvGeom& g, int k, bool fprop) {
    const size_t patch = (((size_t)k * g.PH * g.PW + 7) & ~(size_t)7) * 2;
    return 1024 + 6 * (size_t)kT16 + 64 + 16 + 1024 + 2 * patch + 64;
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
    const size_t smem = in_smem_bytes(g, k, true);
    NIDT_REQUIRE(smem <= 227 * 1024, "conv_in_fprop: patch too large for shared memory (%zu B)", smem);
    const int grid = g.total_tiles < kNumSMs ? g.total_tiles : kNumSMs;
    cudaStream_t s = (cudaStream_t)stream;
    if (k == 5) {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv_in_fprop_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        conv_in_fprop_kernel<5><<<grid, kInThreads, smem, s>>>((const uint8_t*)x, w, bias, ty, stats_partials, g);
    } else {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv_in_fprop_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        conv_in_fprop_kernel<3><<<grid, kInThreads, smem, s>>>((const uint8_t*)x, w, bias, ty, stats_partials, g);
    }
    NIDT_LAUNCH_CHECK();
    if (stats_partials != nullptr && grid < kNumSMs)   // unused partial rows must read as zero
        NIDT_CUDA_CHECK(cudaMemsetAsync(stats_partials + (size_t)grid * 128, 0, (size_t)(kNumSMs - grid) * 128 * sizeof(float), s));
    return 0;
}

// dy: bf16 [N][Do][Ho][Wo][64]; grad: fp32 [64][k^3]; scratch: >= nidt_conv_in_grid()*128*64 floats
NIDT_API int nidt_conv_in_wgrad(const void* x, const void* dy, float* grad, float* scratch, int N, int Di, int Hi, int Wi, int k, int stride, int pad,
                                float scale, void* stream) {
    NIDT_REQUIRE(k == 5 || k == 3, "conv_in_wgrad: filter extent %d unsupported (3 or 5)", k);
    InConvGeom g;
    if (make_geom(g, N, Di, Hi, Wi, k, stride, pad, scale)) return 2;
    CUtensorMap tdy;
    if (make_tmap_act5d(&tdy, dy, N, g.Do, g.Ho, g.Wo, 64, g.Wo, g.TH, 1, 1, 1)) return 1;
    const size_t smem = in_smem_bytes(g, k, false);
    NIDT_REQUIRE(smem <= 227 * 1024, "conv_in_wgrad: patch too large for shared memory (%zu B)", smem);
    const int grid = g.total_tiles < kNumSMs ? g.total_tiles : kNumSMs;
    cudaStream_t s = (cudaStream_t)stream;
    if (k == 5) {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv_in_wgrad_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        conv_in_wgrad_kernel<5><<<grid, kInThreads, smem, s>>>((const uint8_t*)x, tdy, scratch, g);
    } else {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(conv_in_wgrad_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        conv_in_wgrad_kernel<3><<<grid, kInThreads, smem, s>>>((const uint8_t*)x, tdy, scratch, g);
    }
    NIDT_LAUNCH_CHECK();
    const int taps = k * k * k;
    conv_in_wgrad_reduce_kernel<<<(64 * taps + 255) / 256, 256, 0, s>>>(scratch, grid, taps, grad);
    NIDT_LAUNCH_CHECK();
    return 0;
}
