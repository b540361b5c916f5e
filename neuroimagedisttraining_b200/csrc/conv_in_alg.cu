// Weight gradient of the first layer (Conv3d(1,64,5,stride 2) -> BatchNorm3d -> ReLU -> MaxPool3d(3,3)) WITHOUT a pass over the
// layer's full-resolution output: neither y (506 MB at batch 16) nor dY is read or written in the backward pass.
//
// With  dY = s*(dz - c1 - yhat*c2),  s = gamma*invstd,  yhat = (y - mu)*invstd,  dz = the pooled gradient routed to the winning
// position of every pooling window (zero elsewhere), c1 = mean(dz), c2 = mean(dz*yhat)  (bn_pool.cu), and y = W x~ + b linear in the
// input patches x~[p (+) t] (t = one of the 125 taps), the weight gradient splits into
//
//   dW[c][t] = s_c * (  S[c][t]  -  c1_c * sx[t]  -  c2_c * invstd_c * ( (W G)[c][t] + (b_c - mu_c) * sx[t] )  )
//
//   S[c][t]  = sum over pooling windows w of  da[w][c] * x~[p*(w,c) (+) t]      one patch per window and channel: 1/27 of the positions
//   G[t'][t] = sum over ALL output positions p of x~[p (+) t'] * x~[p (+) t]    the Gram matrix of the input patches, 125 x 125
//   sx[t]    = sum over all p of x~[p (+) t]                                    (row 125 of the padded 128 x 128 G: a constant-1 tap)
//
// G and sx depend on the INPUT VOLUME only -- not on the weights, not on the step.  They are computed once per sample
// (gram_plane_kernel + gram_reduce_kernel, ~0.2 ms) and kept on the device keyed by the sample's id (ops/engine_b200.py: GramCache);
// a step adds up the matrices of its 16 samples (gram_sum_kernel).  What is left per step is S (alg_sparse_kernel: 8.5 M patch
// reads instead of 4 M positions x 64 channels x 216 taps on the tensor cores and two 506 MB passes) and a 64 x 125 x 125 fp64
// product (alg_combine_kernel).  The bias gradient is exactly zero (BatchNorm subtracts the batch mean right after the conv).
//
// x~ = bf16(u8 * scale) and W~ = bf16(W) are the operands the forward kernel (conv_in_s2d.cu) multiplies, so this is the gradient of
// the function the forward pass computes; y enters un-rounded, which the dY-based kernels cannot do.
// Reference op: autograd through Conv3d + BatchNorm3d + ReLU + MaxPool3d, fedml_api/model/cv/salient_models.py:147-150 (cuDNN).
#include "common.cuh"
#include "../../include/nidt_b200.h"

namespace nidt {
namespace alg {

constexpr int kTaps = 125;        // 5 x 5 x 5
constexpr int kTP = 128;          // padded: tap 125 = constant 1 (gives sx and the position count), 126 / 127 = 0
constexpr int kCout = 64;
constexpr int kMaxList = 64;

struct IntList { int n; int v[kMaxList]; };

__device__ __forceinline__ float xt(unsigned char b, float scale) { return __bfloat162float(__float2bfloat16((float)b * scale)); }

// ------------------------------------------------------------------------------------------------ per-sample Gram matrix
// grid (Do, samples); 256 threads; thread (ti, tj) owns G[8ti .. 8ti+7][8tj .. 8tj+7] for the positions of ONE output plane
// (Ho x Wo <= ~4.2 k positions: short fp32 sums), written to partials[sample][plane][128][128]; gram_reduce adds the planes in fp64.
// Shared memory per output line: the 5 x 5 input rows it reads (as x~), a row of ones, a row of zeros; a tap is a row + a column
// offset, an output position adds 2*w to the column.
__global__ void __launch_bounds__(256) gram_plane_kernel(const unsigned char* __restrict__ x, const IntList rows, int Di, int Hi, int Wi, int Ho, int Wo,
                                                         float scale, float* __restrict__ partials) {
    extern __shared__ float sm[];
    const int pitch = Wi + 7;                       // >= 2*(Wo-1) + 4 + 1
    float* ones = sm + 25 * pitch;
    float* zeros = ones + pitch;
    const int d = blockIdx.x, sample = blockIdx.y;
    const int Do = gridDim.x;
    const unsigned char* vol = x + (size_t)rows.v[sample] * Di * Hi * Wi;
    const int tid = threadIdx.x, ti = tid >> 4, tj = tid & 15;
    for (int i = tid; i < pitch; i += 256) { ones[i] = 1.f; zeros[i] = 0.f; }
    int offi[8], offj[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int a = ti * 8 + k, b = tj * 8 + k;
        offi[k] = a < kTaps ? ((a / 25) * 5 + (a / 5) % 5) * pitch + a % 5 : (a == kTaps ? 25 * pitch : 26 * pitch);
        offj[k] = b < kTaps ? ((b / 25) * 5 + (b / 5) % 5) * pitch + b % 5 : (b == kTaps ? 25 * pitch : 26 * pitch);
    }
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    for (int h = 0; h < Ho; ++h) {
        __syncthreads();
        for (int i = tid; i < 25 * Wi; i += 256) {
            const int r = i / Wi, col = i % Wi;
            sm[r * pitch + col] = xt(vol[((size_t)(2 * d + r / 5) * Hi + (2 * h + r % 5)) * Wi + col], scale);
        }
        __syncthreads();
        for (int w = 0; w < Wo; ++w) {
            float vi[8], vj[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) { vi[k] = sm[offi[k] + 2 * w]; vj[k] = sm[offj[k] + 2 * w]; }
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(vi[i], vj[j], acc[i][j]);
        }
    }
    float* out = partials + ((size_t)sample * Do + d) * (kTP * kTP);
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; j += 4)
            *reinterpret_cast<float4*>(out + (ti * 8 + i) * kTP + tj * 8 + j) = make_float4(acc[i][j], acc[i][j + 1], acc[i][j + 2], acc[i][j + 3]);
}

// cache[slot[sample]][e] = sum over planes of partials[sample][plane][e]   (fp64 sum, fp32 result)
__global__ void __launch_bounds__(256) gram_reduce_kernel(const float* __restrict__ partials, int planes, const IntList slots, float* __restrict__ cache) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    const int sample = blockIdx.y;
    const float* p = partials + (size_t)sample * planes * (kTP * kTP) + e;
    double a = 0.0;
    for (int d = 0; d < planes; ++d) a += (double)p[(size_t)d * (kTP * kTP)];
    cache[(size_t)slots.v[sample] * (kTP * kTP) + e] = (float)a;
}

// gb[e] = sum over the batch's samples of cache[slot][e]   (fp64)
__global__ void __launch_bounds__(256) gram_sum_kernel(const float* __restrict__ cache, const IntList slots, double* __restrict__ gb) {
    const int e = blockIdx.x * 256 + threadIdx.x;
    double a = 0.0;
    for (int k = 0; k < slots.n; ++k) a += (double)cache[(size_t)slots.v[k] * (kTP * kTP) + e];
    gb[e] = a;
}

// ------------------------------------------------------------------------------------------------ sparse term
// One CTA walks pooling windows (n, dp, hp, wp).  Per window: the 9 x 9 x 9 input neighbourhood it covers goes to shared memory
// as x~ (double-buffered), then thread (c = tid % 64, tg = tid / 64) adds da[w][c] * x~[winner patch of (w,c)][t] for its 32 taps
// t = 32*tg ..  into registers.  partials[cta][c][128].
constexpr int kNbPitch = 9;
constexpr int kNbPlane = 9 * kNbPitch;
constexpr int kNbSize = 9 * kNbPlane;       // 729

__global__ void __launch_bounds__(256) alg_sparse_kernel(const unsigned char* __restrict__ x, const unsigned char* __restrict__ amax,
                                                        const __nv_bfloat16* __restrict__ da, int N, int Di, int Hi, int Wi, int Dp, int Hp, int Wp,
                                                        float scale, float* __restrict__ partials) {
    __shared__ float nb[2][kNbSize + 3];
    const int tid = threadIdx.x, c = tid & 63, tg = tid >> 6;
    int toff[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) {
        const int t = tg * 32 + k;
        toff[k] = t < kTaps ? (t / 25) * kNbPlane + ((t / 5) % 5) * kNbPitch + t % 5 : -1;
    }
    float acc[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) acc[k] = 0.f;
    const int total = N * Dp * Hp * Wp;
    auto load = [&](int w, int buf) {
        const int wp = w % Wp, hp = (w / Wp) % Hp, dp = (w / (Wp * Hp)) % Dp, n = w / (Wp * Hp * Dp);
        const unsigned char* base = x + (((size_t)n * Di + 6 * dp) * Hi + 6 * hp) * Wi + 6 * wp;
        for (int i = tid; i < kNbSize; i += 256) {
            const int z = i / kNbPlane, r = (i / kNbPitch) % 9, col = i % kNbPitch;
            nb[buf][i] = xt(base[((size_t)z * Hi + r) * Wi + col], scale);
        }
    };
    int w = blockIdx.x, it = 0;
    if (w < total) load(w, 0);
    for (; w < total; w += gridDim.x, ++it) {
        __syncthreads();                                   // nb[it & 1] is complete; everybody is done with nb[(it + 1) & 1]
        const int wn = w + gridDim.x;
        if (wn < total) load(wn, (it + 1) & 1);
        const unsigned int am = amax[(size_t)w * kCout + c];
        if (am != 255u) {
            const float g = __bfloat162float(da[(size_t)w * kCout + c]);
            const float* p = nb[it & 1] + (2 * (am / 9)) * kNbPlane + (2 * ((am / 3) % 3)) * kNbPitch + 2 * (am % 3);
#pragma unroll
            for (int k = 0; k < 32; ++k)
                if (toff[k] >= 0) acc[k] = fmaf(g, p[toff[k]], acc[k]);
        }
    }
    float* out = partials + ((size_t)blockIdx.x * kCout + c) * kTP + tg * 32;
#pragma unroll
    for (int k = 0; k < 32; k += 4) *reinterpret_cast<float4*>(out + k) = make_float4(acc[k], acc[k + 1], acc[k + 2], acc[k + 3]);
}

// ------------------------------------------------------------------------------------------------ combine
// grid 64 (channel), 128 threads (tap).  dW as in the file header; fp64 throughout (the bracket is a difference of sums over 4 M
// positions).  Optional SNIP epilogue: scores += sal_scale * |dW * w|.
__global__ void __launch_bounds__(128) alg_combine_kernel(const float* __restrict__ partials, int n_part, const double* __restrict__ gb,
                                                         const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ bn_scale,
                                                         const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ c12,
                                                         float* __restrict__ grad_w, float* __restrict__ grad_b, const float* __restrict__ sal_w,
                                                         float* __restrict__ sal_scores, float sal_scale) {
    __shared__ double wq[kTP];
    const int c = blockIdx.x, t = threadIdx.x;
    wq[t] = t < kTaps ? (double)__bfloat162float(__float2bfloat16(W[c * kTaps + t])) : 0.0;
    __syncthreads();
    double S = 0.0;
    for (int b = 0; b < n_part; ++b) S += (double)partials[((size_t)b * kCout + c) * kTP + t];
    double wg = 0.0;
    for (int u = 0; u < kTaps; ++u) wg = fma(wq[u], gb[u * kTP + t], wg);
    const double sx = gb[kTaps * kTP + t];
    const double is = (double)invstd[c], mu = (double)mean[c], k1 = (double)c12[c], k2 = (double)c12[kCout + c];
    const double bc = bias != nullptr ? (double)bias[c] : 0.0;
    const double dw = (double)bn_scale[c] * (S - k1 * sx - k2 * is * (wg + (bc - mu) * sx));
    if (t < kTaps) {
        const int o = c * kTaps + t;
        grad_w[o] = (float)dw;
        if (sal_scores != nullptr) sal_scores[o] = fmaf(sal_scale, fabsf((float)dw * sal_w[o]), sal_scores[o]);
    }
    if (t == 0 && grad_b != nullptr) grad_b[c] = 0.f;
}

}  // namespace alg
}  // namespace nidt
using namespace nidt;

static bool alg_shape_ok(int Di, int Hi, int Wi) { return Di >= 5 && Hi >= 5 && Wi >= 5 && Wi <= 512; }

NIDT_API int nidt_conv_in_gram_floats(void) { return alg::kTP * alg::kTP; }
NIDT_API int nidt_conv_in_gram_max_samples(void) { return alg::kMaxList; }
NIDT_API int nidt_conv_in_alg_grid(void) { return 4 * kNumSMs; }
// floats of scratch for nidt_conv_in_gram_samples over n samples of depth Di
NIDT_API int64_t nidt_conv_in_gram_scratch_floats(int n, int Di) { return (int64_t)n * ((Di - 5) / 2 + 1) * alg::kTP * alg::kTP; }
// floats of scratch for nidt_conv_in_alg_wgrad
NIDT_API int64_t nidt_conv_in_alg_scratch_floats(void) { return (int64_t)nidt_conv_in_alg_grid() * alg::kCout * alg::kTP; }

// Gram matrices (128 x 128 fp32, tap 125 = constant 1) of n samples: sample k is row rows[k] of the u8 batch x [*, Di, Hi, Wi] and
// goes to cache + slots[k] * 16384.  rows / slots are HOST arrays (n <= 64); scratch: nidt_conv_in_gram_scratch_floats(n, Di).
NIDT_API int nidt_conv_in_gram_samples(const void* x, const int* rows, const int* slots, int n, float* cache, float* scratch, int Di, int Hi,
                                       int Wi, float scale, void* stream) {
    NIDT_REQUIRE(n >= 0 && n <= alg::kMaxList, "nidt_conv_in_gram_samples: n=%d (max %d)", n, alg::kMaxList);
    NIDT_REQUIRE(alg_shape_ok(Di, Hi, Wi), "nidt_conv_in_gram_samples: unsupported volume %dx%dx%d", Di, Hi, Wi);
    if (n == 0) return 0;
    alg::IntList r{}, s{};
    r.n = s.n = n;
    for (int i = 0; i < n; ++i) { r.v[i] = rows[i]; s.v[i] = slots[i]; }
    const int Do = (Di - 5) / 2 + 1, Ho = (Hi - 5) / 2 + 1, Wo = (Wi - 5) / 2 + 1;
    const size_t smem = (size_t)27 * (Wi + 7) * sizeof(float);
    cudaStream_t st = (cudaStream_t)stream;
    alg::gram_plane_kernel<<<dim3(Do, n), 256, smem, st>>>((const unsigned char*)x, r, Di, Hi, Wi, Ho, Wo, scale, scratch);
    NIDT_LAUNCH_CHECK();
    alg::gram_reduce_kernel<<<dim3(alg::kTP * alg::kTP / 256, n), 256, 0, st>>>(scratch, Do, s, cache);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// gb (fp64 [128][128]) = sum of the cached matrices in slots[0..n)   (HOST array, n <= 64)
NIDT_API int nidt_conv_in_gram_sum(const float* cache, const int* slots, int n, double* gb, void* stream) {
    NIDT_REQUIRE(n >= 1 && n <= alg::kMaxList, "nidt_conv_in_gram_sum: n=%d (max %d)", n, alg::kMaxList);
    alg::IntList s{};
    s.n = n;
    for (int i = 0; i < n; ++i) s.v[i] = slots[i];
    alg::gram_sum_kernel<<<alg::kTP * alg::kTP / 256, 256, 0, (cudaStream_t)stream>>>(cache, s, gb);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// dW (and a zero bias gradient) of the first layer from the pooled tensors; see the file header.  x: u8 [N][Di][Hi][Wi]; amax / da:
// [N][Dp][Hp][Wp][64] (winning taps from nidt_bn_relu_pool3_fwd_amax, gradient w.r.t. the pooled activation); W / bias: the fp32 conv
// parameters; bn_scale / mean / invstd / c12: BatchNorm forward scale, batch statistics and backward means (nidt_bn_relu_pool3_bwd_stats);
// gb: the batch's Gram matrix (nidt_conv_in_gram_sum).  scratch: nidt_conv_in_alg_scratch_floats().
NIDT_API int nidt_conv_in_alg_wgrad(const void* x, const void* amax, const void* da, const float* W, const float* bias, const float* bn_scale,
                                    const float* mean, const float* invstd, const float* c12, const double* gb, float* grad_w, float* grad_b,
                                    float* scratch, int N, int Di, int Hi, int Wi, float scale, const float* sal_w, float* sal_scores,
                                    float sal_scale, void* stream) {
    NIDT_REQUIRE(alg_shape_ok(Di, Hi, Wi), "nidt_conv_in_alg_wgrad: unsupported volume %dx%dx%d", Di, Hi, Wi);
    NIDT_REQUIRE(x && amax && da && W && bn_scale && mean && invstd && c12 && gb && grad_w && scratch, "nidt_conv_in_alg_wgrad: null argument");
    const int Do = (Di - 5) / 2 + 1, Ho = (Hi - 5) / 2 + 1, Wo = (Wi - 5) / 2 + 1;
    const int Dp = (Do - 3) / 3 + 1, Hp = (Ho - 3) / 3 + 1, Wp = (Wo - 3) / 3 + 1;
    NIDT_REQUIRE(Do >= 3 && Ho >= 3 && Wo >= 3, "nidt_conv_in_alg_wgrad: conv output %dx%dx%d smaller than the pooling window", Do, Ho, Wo);
    const int total = N * Dp * Hp * Wp;
    int grid = nidt_conv_in_alg_grid();
    if (grid > total) grid = total;
    cudaStream_t st = (cudaStream_t)stream;
    alg::alg_sparse_kernel<<<grid, 256, 0, st>>>((const unsigned char*)x, (const unsigned char*)amax, (const __nv_bfloat16*)da, N, Di, Hi, Wi, Dp,
                                                 Hp, Wp, scale, scratch);
    NIDT_LAUNCH_CHECK();
    alg::alg_combine_kernel<<<alg::kCout, 128, 0, st>>>(scratch, grid, gb, W, bias, bn_scale, mean, invstd, c12, grad_w, grad_b, sal_w, sal_scores,
                                                       sal_scale);
    NIDT_LAUNCH_CHECK();
    return 0;
}
