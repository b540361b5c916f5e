// Fused BatchNorm3d (training or eval statistics) + ReLU + MaxPool3d on channels-last (NDHWC) bf16
// activations, forward and backward.  Replaces K3+K4+K5 of SURVEY §2.5 (cuDNN BN, ATen ReLU,
// ATen MaxPool3d fwd/bwd = 6 library kernels and 4 full-size intermediates per block in the
// reference, fedml_api/model/cv/salient_models.py:147-168) with:
//
//   conv epilogue        -> per-tile (sum, sum^2) partials               (conv kernels)
//   nidt_bn_finalize     -> mean / invstd / scale / shift, running stats  (tiny)
//   nidt_bn_relu_pool_fwd-> a = maxpool(relu(y*scale+shift))              (1 read of y, 1/27 write)
//   nidt_bn_relu_pool_bwd_reduce -> sum(dyhat), sum(dyhat*xhat) partials  (argmax recomputed from y)
//   nidt_bn_bwd_finalize -> c1, c2, dgamma, dbeta
//   nidt_bn_relu_pool_bwd_apply  -> dy = scale*(dyhat - c1 - xhat*c2)     (input of dgrad / wgrad)
//
// No argmax indices, no normalised activations and no ReLU masks are ever stored.
#include "common.cuh"
#include <stdlib.h>

namespace nidt {

struct BlockGeom {
    int N, D, H, W, C;       // conv output (pre-pool) extents
    int pk, ps;              // pool kernel / stride (0 = no pooling)
    int Dp, Hp, Wp;          // pooled extents (== D,H,W when pk == 0)
};

__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(p[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
    uint4 v;
    __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return v;
}

// ------------------------------------------------------------------------------- finalize
// partials: [T][2][C] (sum, sumsq) per conv tile.  One block per 16 channels, 64 tile-slices (T reaches a few thousand
// for the wide early layers; double accumulation keeps the result independent of the slicing to fp32 round-off).
constexpr int kFinCh = 16, kFinSlices = 64;
__global__ void __launch_bounds__(kFinCh * kFinSlices) bn_finalize_kernel(
    const float* __restrict__ partials, int T, int C, double count, const float* __restrict__ gamma, const float* __restrict__ beta,
    float* __restrict__ running_mean, float* __restrict__ running_var, float* __restrict__ num_batches, float momentum, float eps,
    float* __restrict__ scale, float* __restrict__ shift, float* __restrict__ mean_out, float* __restrict__ invstd_out) {
    __shared__ double s1[kFinSlices][kFinCh], s2[kFinSlices][kFinCh];
    const int tx = threadIdx.x % kFinCh, ty = threadIdx.x / kFinCh;
    const int c = blockIdx.x * kFinCh + tx;
    double a = 0.0, b = 0.0;
    if (c < C) {
        int t = ty;
        for (; t + 3 * kFinSlices < T; t += 4 * kFinSlices) {      // 8 independent loads in flight
            float v[8];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                v[2 * u] = partials[((size_t)(t + u * kFinSlices) * 2 + 0) * C + c];
                v[2 * u + 1] = partials[((size_t)(t + u * kFinSlices) * 2 + 1) * C + c];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) { a += (double)v[2 * u]; b += (double)v[2 * u + 1]; }
        }
        for (; t < T; t += kFinSlices) {
            a += (double)partials[((size_t)t * 2 + 0) * C + c];
            b += (double)partials[((size_t)t * 2 + 1) * C + c];
        }
    }
    s1[ty][tx] = a;
    s2[ty][tx] = b;
    __syncthreads();
    if (ty == 0 && c < C) {
        for (int i = 1; i < kFinSlices; ++i) { a += s1[i][tx]; b += s2[i][tx]; }
        const double mean = a / count;
        double var = b / count - mean * mean;
        if (var < 0.0) var = 0.0;
        const float invstd = (float)(1.0 / sqrt(var + (double)eps));
        const float g = gamma[c];
        scale[c] = g * invstd;
        shift[c] = beta[c] - (float)mean * g * invstd;
        mean_out[c] = (float)mean;
        invstd_out[c] = invstd;
        if (running_mean != nullptr) {
            const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
            running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches != nullptr) *num_batches += 1.f;
}

// eval mode: scale/shift from the running statistics
__global__ void bn_eval_prepare_kernel(int C, const float* __restrict__ gamma, const float* __restrict__ beta,
                                       const float* __restrict__ running_mean, const float* __restrict__ running_var, float eps,
                                       float* __restrict__ scale, float* __restrict__ shift) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) {
        const float s = gamma[c] * rsqrtf(running_var[c] + eps);
        scale[c] = s;
        shift[c] = beta[c] - running_mean[c] * s;
    }
}

// -------------------------------------------------------------------------------- forward
// one thread = one output position x 8 channels
__global__ void __launch_bounds__(256) bn_relu_pool_fwd_kernel(
    const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ a, BlockGeom g,
    const float* __restrict__ scale, const float* __restrict__ shift, int relu) {
    const int CG = g.C >> 3;
    const int64_t total = (int64_t)g.N * g.Dp * g.Hp * g.Wp * CG;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int cg = (int)(idx % CG);
        int64_t p = idx / CG;
        const int wp = (int)(p % g.Wp); p /= g.Wp;
        const int hp = (int)(p % g.Hp); p /= g.Hp;
        const int dp = (int)(p % g.Dp);
        const int n = (int)(p / g.Dp);
        float sc[8], sh[8], best[8];
        *reinterpret_cast<float4*>(sc) = ld4(scale + cg * 8);
        *reinterpret_cast<float4*>(sc + 4) = ld4(scale + cg * 8 + 4);
        *reinterpret_cast<float4*>(sh) = ld4(shift + cg * 8);
        *reinterpret_cast<float4*>(sh + 4) = ld4(shift + cg * 8 + 4);
        if (g.pk == 0) {
            const uint4 v = *reinterpret_cast<const uint4*>(y + ((((int64_t)n * g.D + dp) * g.H + hp) * g.W + wp) * g.C + cg * 8);
            unpack8(v, best);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                best[i] = fmaf(best[i], sc[i], sh[i]);
                if (relu) best[i] = fmaxf(best[i], 0.f);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) best[i] = -INFINITY;
            for (int kd = 0; kd < g.pk; ++kd)
                for (int kh = 0; kh < g.pk; ++kh) {
                    const __nv_bfloat16* row = y + ((((int64_t)n * g.D + dp * g.ps + kd) * g.H + hp * g.ps + kh) * g.W + wp * g.ps) * g.C + cg * 8;
                    for (int kw = 0; kw < g.pk; ++kw) {
                        float f[8];
                        unpack8(*reinterpret_cast<const uint4*>(row + (int64_t)kw * g.C), f);
#pragma unroll
                        for (int i = 0; i < 8; ++i) best[i] = fmaxf(best[i], fmaf(f[i], sc[i], sh[i]));
                    }
                }
            if (relu) {
#pragma unroll
                for (int i = 0; i < 8; ++i) best[i] = fmaxf(best[i], 0.f);
            }
        }
        *reinterpret_cast<uint4*>(a + idx * 8) = pack8(best);
    }
}

// ------------------------------------------------------------------------------- backward
// Shared by reduce (write_dy = 0) and apply (write_dy = 1).  One thread = one pooling window
// (or one position when pk == 0) x 8 channels.  Windows are enumerated over ceil(D/ps) etc. so
// that the remainder positions no pooled output covers are visited too (their dyhat is 0).
template <bool kApply>
__global__ void __launch_bounds__(256) bn_relu_pool_bwd_kernel(
    const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ da, __nv_bfloat16* __restrict__ dy, BlockGeom g,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean, const float* __restrict__ invstd,
    const float* __restrict__ c12 /* [2][C]: c1, c2 */, float* __restrict__ partials /* [grid][2][C] */, int relu) {
    extern __shared__ float sh_red[];   // reduce only: [2][PB][C]
    const int CG = g.C >> 3;
    const int PB = blockDim.x / CG;     // positions per block pass (blockDim.x == PB*CG)
    const int cg = threadIdx.x % CG;
    const int pl = threadIdx.x / CG;
    const int step = g.pk ? g.ps : 1;
    const int Dw = g.pk ? (g.D + step - 1) / step : g.D;
    const int Hw = g.pk ? (g.H + step - 1) / step : g.H;
    const int Ww = g.pk ? (g.W + step - 1) / step : g.W;
    const int64_t nwin = (int64_t)g.N * Dw * Hw * Ww;
    float sc[8], shf[8], mu[8], is[8], k1[8], k2[8], acc1[8], acc2[8];
    *reinterpret_cast<float4*>(sc) = ld4(scale + cg * 8);   *reinterpret_cast<float4*>(sc + 4) = ld4(scale + cg * 8 + 4);
    *reinterpret_cast<float4*>(shf) = ld4(shift + cg * 8);  *reinterpret_cast<float4*>(shf + 4) = ld4(shift + cg * 8 + 4);
    *reinterpret_cast<float4*>(mu) = ld4(mean + cg * 8);    *reinterpret_cast<float4*>(mu + 4) = ld4(mean + cg * 8 + 4);
    *reinterpret_cast<float4*>(is) = ld4(invstd + cg * 8);  *reinterpret_cast<float4*>(is + 4) = ld4(invstd + cg * 8 + 4);
    if (kApply) {
        *reinterpret_cast<float4*>(k1) = ld4(c12 + cg * 8);        *reinterpret_cast<float4*>(k1 + 4) = ld4(c12 + cg * 8 + 4);
        *reinterpret_cast<float4*>(k2) = ld4(c12 + g.C + cg * 8);  *reinterpret_cast<float4*>(k2 + 4) = ld4(c12 + g.C + cg * 8 + 4);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc1[i] = acc2[i] = 0.f;

    for (int64_t win = (int64_t)blockIdx.x * PB + pl; win < nwin; win += (int64_t)gridDim.x * PB) {
        if (pl >= PB) break;
        int64_t p = win;
        const int ww = (int)(p % Ww); p /= Ww;
        const int hw = (int)(p % Hw); p /= Hw;
        const int dw = (int)(p % Dw);
        const int n = (int)(p / Dw);
        if (g.pk == 0) {
            const int64_t off = ((((int64_t)n * g.D + dw) * g.H + hw) * g.W + ww) * g.C + cg * 8;
            float f[8], d[8];
            unpack8(*reinterpret_cast<const uint4*>(y + off), f);
            unpack8(*reinterpret_cast<const uint4*>(da + off), d);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float xh = (f[i] - mu[i]) * is[i];
                const float dyh = (!relu || fmaf(f[i], sc[i], shf[i]) > 0.f) ? d[i] : 0.f;
                if (kApply) d[i] = sc[i] * (dyh - k1[i] - xh * k2[i]);
                else { acc1[i] += dyh; acc2[i] = fmaf(dyh, xh, acc2[i]); }
            }
            if (kApply) *reinterpret_cast<uint4*>(dy + off) = pack8(d);
            continue;
        }
        const bool has_out = dw < g.Dp && hw < g.Hp && ww < g.Wp;
        float best[8], dav[8];
        int arg[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { best[i] = -INFINITY; arg[i] = -1; dav[i] = 0.f; }
        if (has_out) {
            unpack8(*reinterpret_cast<const uint4*>(da + ((((int64_t)n * g.Dp + dw) * g.Hp + hw) * g.Wp + ww) * g.C + cg * 8), dav);
            for (int kd = 0; kd < g.pk; ++kd)
                for (int kh = 0; kh < g.pk; ++kh) {
                    const __nv_bfloat16* row = y + ((((int64_t)n * g.D + dw * step + kd) * g.H + hw * step + kh) * g.W + ww * step) * g.C + cg * 8;
                    for (int kw = 0; kw < g.pk; ++kw) {
                        float f[8];
                        unpack8(*reinterpret_cast<const uint4*>(row + (int64_t)kw * g.C), f);
                        const int t = (kd * g.pk + kh) * g.pk + kw;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float v = fmaf(f[i], sc[i], shf[i]);   // pooling runs on the post-BN value
                            if (v > best[i]) { best[i] = v; arg[i] = t; }
                        }
                    }
                }
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (relu && !(best[i] > 0.f)) arg[i] = -1;             // ReLU gate on the winner
        }
        if (!kApply) {
            if (has_out) {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (arg[i] >= 0) {
                        // xhat of the winner, reconstructed from its post-BN value: v = gamma*xhat + beta
                        // is avoided (gamma may be 0); re-read the winner's y instead.
                        const int t = arg[i];
                        const int kd = t / (g.pk * g.pk), kh = (t / g.pk) % g.pk, kw = t % g.pk;
                        const float yv = __bfloat162float(
                            y[((((int64_t)n * g.D + dw * step + kd) * g.H + hw * step + kh) * g.W + ww * step + kw) * g.C + cg * 8 + i]);
                        const float xh = (yv - mu[i]) * is[i];
                        acc1[i] += dav[i];
                        acc2[i] = fmaf(dav[i], xh, acc2[i]);
                    }
            }
        } else {
            for (int kd = 0; kd < g.pk; ++kd) {
                const int d_ = dw * step + kd;
                if (d_ >= g.D) break;
                for (int kh = 0; kh < g.pk; ++kh) {
                    const int h_ = hw * step + kh;
                    if (h_ >= g.H) break;
                    for (int kw = 0; kw < g.pk; ++kw) {
                        const int w_ = ww * step + kw;
                        if (w_ >= g.W) break;
                        const int64_t off = ((((int64_t)n * g.D + d_) * g.H + h_) * g.W + w_) * g.C + cg * 8;
                        float f[8], o[8];
                        unpack8(*reinterpret_cast<const uint4*>(y + off), f);
                        const int t = (kd * g.pk + kh) * g.pk + kw;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const float xh = (f[i] - mu[i]) * is[i];
                            const float dyh = (arg[i] == t) ? dav[i] : 0.f;
                            o[i] = sc[i] * (dyh - k1[i] - xh * k2[i]);
                        }
                        *reinterpret_cast<uint4*>(dy + off) = pack8(o);
                    }
                }
            }
        }
    }
    if (!kApply) {
        // deterministic block reduction over the PB positions that share a channel group
        float* r1 = sh_red;
        float* r2 = sh_red + PB * g.C;
        if (pl < PB) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                r1[pl * g.C + cg * 8 + i] = acc1[i];
                r2[pl * g.C + cg * 8 + i] = acc2[i];
            }
        }
        __syncthreads();
        for (int c = threadIdx.x; c < g.C; c += blockDim.x) {
            float a = 0.f, b = 0.f;
            for (int q = 0; q < PB; ++q) { a += r1[q * g.C + c]; b += r2[q * g.C + c]; }
            partials[((size_t)blockIdx.x * 2 + 0) * g.C + c] = a;
            partials[((size_t)blockIdx.x * 2 + 1) * g.C + c] = b;
        }
    }
}

// ---------------------------------------------------------------- 3x3x3 / stride-3 fast paths
// Forward: also emits ymax = the raw conv output at the window's arg-max (first maximum in d,h,w scan
// order, PyTorch's rule).  With it the backward needs no arg-max search and no second pass over y:
//   reduce: elementwise over the *pooled* tensor (1/27 of the data),
//   apply : one streaming pass; the winner is the first element whose post-BN value equals the
//           post-BN value of ymax.
template <int V> struct VecT;
template <> struct VecT<8> { using T = uint4; };
template <> struct VecT<4> { using T = uint2; };
template <int V>
__device__ __forceinline__ void unpackv(const typename VecT<V>::T& v, float* f) {
    const __nv_bfloat162* p = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < V / 2; ++i) {
        float2 t = __bfloat1622float2(p[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
template <int V>
__device__ __forceinline__ typename VecT<V>::T packv(const float* f) {
    typename VecT<V>::T v;
    __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < V / 2; ++i) p[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return v;
}
template <int V>
__device__ __forceinline__ void ldv(const float* p, float* f) {
#pragma unroll
    for (int i = 0; i < V; i += 4) *reinterpret_cast<float4*>(f + i) = ld4(p + i);
}

// one thread = one pooled output x V channels (V = 4: twice the threads, half the registers -> more loads in flight)
template <int V, bool kAmax>
__global__ void __launch_bounds__(256, V == 4 ? 3 : 1) bn_relu_pool3_fwd_kernel(
    const __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ a, __nv_bfloat16* __restrict__ ymax, BlockGeom g,
    const float* __restrict__ scale, const float* __restrict__ shift, int relu, unsigned char* __restrict__ amax) {
    using VT = typename VecT<V>::T;
    const int CG = g.C / V;
    const int64_t total = (int64_t)g.N * g.Dp * g.Hp * g.Wp * CG;
    unsigned char ab[V];
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int cg = (int)(idx % CG);
    int64_t p = idx / CG;
    const int wp = (int)(p % g.Wp); p /= g.Wp;
    const int hp = (int)(p % g.Hp); p /= g.Hp;
    const int dp = (int)(p % g.Dp);
    const int n = (int)(p / g.Dp);
    float sc[V], sh[V], best[V], yb[V];
    ldv<V>(scale + cg * V, sc);
    ldv<V>(shift + cg * V, sh);
#pragma unroll
    for (int i = 0; i < V; ++i) { best[i] = -INFINITY; yb[i] = 0.f; ab[i] = 0; }
    const __nv_bfloat16* base = y + ((((int64_t)n * g.D + dp * 3) * g.H + hp * 3) * g.W + wp * 3) * g.C + cg * V;
    const int64_t sH = (int64_t)g.W * g.C, sD = (int64_t)g.H * g.W * g.C;
#pragma unroll
    for (int kd = 0; kd < 3; ++kd) {
        VT v[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) v[t] = *reinterpret_cast<const VT*>(base + kd * sD + (t / 3) * sH + (t % 3) * g.C);
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            float f[V];
            unpackv<V>(v[t], f);
#pragma unroll
            for (int i = 0; i < V; ++i) {
                const float pv = fmaf(f[i], sc[i], sh[i]);
                if (pv > best[i]) {
                    best[i] = pv; yb[i] = f[i];
                    if (kAmax) ab[i] = (unsigned char)(kd * 9 + t);
                }
            }
        }
    }
    if (kAmax) {
        // tap (kd*9 + kh*3 + kw) of the FIRST maximum of every window, 255 when ReLU gates the window (nobody gets a gradient):
        // lets the backward decide "am I the winner" per element without looking at its neighbours (fused wgrad of block 0)
#pragma unroll
        for (int i = 0; i < V; ++i)
            if (relu && !(best[i] > 0.f)) ab[i] = 255;
        // one V-byte store per thread (V = 8: uint2, V = 4: unsigned)
        unsigned int w0 = 0u, w1 = 0u;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            if (i < 4) w0 |= (unsigned int)ab[i] << (8 * i);
            else w1 |= (unsigned int)ab[i] << (8 * (i - 4));
        }
        if (V == 8) *reinterpret_cast<uint2*>(amax + idx * V) = make_uint2(w0, w1);
        else *reinterpret_cast<unsigned int*>(amax + idx * V) = w0;
    }
    if (relu) {
#pragma unroll
        for (int i = 0; i < V; ++i) best[i] = fmaxf(best[i], 0.f);
    }
    *reinterpret_cast<VT*>(a + idx * V) = packv<V>(best);
    *reinterpret_cast<VT*>(ymax + idx * V) = packv<V>(yb);
}

// The same pooling from a conv output that was already reduced along W in the conv epilogue (nidt_conv_in_s2d_fprop_poolw): per window
// 9 candidates (kd, kh) -- each the winner of its 3 columns, with the column in 2 bits -- instead of 27.  One thread = one pooled output
// x 8 channels.  Candidates are visited in (kd, kh) order with a strict comparison, and the column winner was the first maximum of its
// three, so the tap recorded is the first maximum in (kd, kh, kw) order: exactly what bn_relu_pool3_fwd_kernel finds on the full tensor.
template <bool kTrain>
__global__ void __launch_bounds__(256) bn_relu_pool3w_fwd_kernel(const __nv_bfloat16* __restrict__ yw, const unsigned short* __restrict__ aw,
                                                                 __nv_bfloat16* __restrict__ a, __nv_bfloat16* __restrict__ ymax,
                                                                 unsigned char* __restrict__ amax, int N, int D, int H, int Dp, int Hp, int Wp,
                                                                 const float* __restrict__ scale, const float* __restrict__ shift, int relu) {
    constexpr int C = 64, CG = 8;
    const int64_t total = (int64_t)N * Dp * Hp * Wp * CG;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int cg = (int)(idx % CG);
    int64_t p = idx / CG;
    const int wp = (int)(p % Wp); p /= Wp;
    const int hp = (int)(p % Hp); p /= Hp;
    const int dp = (int)(p % Dp);
    const int n = (int)(p / Dp);
    float sc[8], sh[8], best[8], yb[8];
    unsigned int tap[8];
    ldv<8>(scale + cg * 8, sc);
    ldv<8>(shift + cg * 8, sh);
#pragma unroll
    for (int i = 0; i < 8; ++i) { best[i] = -INFINITY; yb[i] = 0.f; tap[i] = 0u; }
    uint4 v[9];
    unsigned short cw[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        const int64_t pos = (((int64_t)n * D + dp * 3 + t / 3) * H + hp * 3 + t % 3) * Wp + wp;
        v[t] = *reinterpret_cast<const uint4*>(yw + pos * C + cg * 8);
        cw[t] = aw[pos * CG + cg];
    }
#pragma unroll
    for (int t = 0; t < 9; ++t) {
        float f[8];
        unpack8(v[t], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float pv = fmaf(f[i], sc[i], sh[i]);
            if (pv > best[i]) {
                best[i] = pv; yb[i] = f[i];
                tap[i] = (unsigned int)((t / 3) * 9 + (t % 3) * 3) + (((unsigned int)cw[t] >> (2 * i)) & 3u);
            }
        }
    }
    if (kTrain) {
        unsigned int w0 = 0u, w1 = 0u;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const unsigned int tb = (relu && !(best[i] > 0.f)) ? 255u : tap[i];
            if (i < 4) w0 |= tb << (8 * i);
            else w1 |= tb << (8 * (i - 4));
        }
        *reinterpret_cast<uint2*>(amax + idx * 8) = make_uint2(w0, w1);
        *reinterpret_cast<uint4*>(ymax + idx * 8) = packv<8>(yb);
    }
    if (relu) {
#pragma unroll
        for (int i = 0; i < 8; ++i) best[i] = fmaxf(best[i], 0.f);
    }
    *reinterpret_cast<uint4*>(a + idx * 8) = packv<8>(best);
}

// reduce over the pooled tensor: partials[block][2][C]
__global__ void __launch_bounds__(256) bn_pool3_bwd_reduce_kernel(
    const __nv_bfloat16* __restrict__ ymax, const __nv_bfloat16* __restrict__ da, int64_t npos, int C,
    const float* __restrict__ scale, const float* __restrict__ shift, const float* __restrict__ mean, const float* __restrict__ invstd,
    float* __restrict__ partials, int relu) {
    extern __shared__ float sh_red[];
    const int CG = C >> 3;
    const int PB = blockDim.x / CG;
    const int cg = threadIdx.x % CG, pl = threadIdx.x / CG;
    float sc[8], shf[8], mu[8], is[8], acc1[8], acc2[8];
    *reinterpret_cast<float4*>(sc) = ld4(scale + cg * 8);   *reinterpret_cast<float4*>(sc + 4) = ld4(scale + cg * 8 + 4);
    *reinterpret_cast<float4*>(shf) = ld4(shift + cg * 8);  *reinterpret_cast<float4*>(shf + 4) = ld4(shift + cg * 8 + 4);
    *reinterpret_cast<float4*>(mu) = ld4(mean + cg * 8);    *reinterpret_cast<float4*>(mu + 4) = ld4(mean + cg * 8 + 4);
    *reinterpret_cast<float4*>(is) = ld4(invstd + cg * 8);  *reinterpret_cast<float4*>(is + 4) = ld4(invstd + cg * 8 + 4);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc1[i] = acc2[i] = 0.f;
    for (int64_t pos = (int64_t)blockIdx.x * PB + pl; pos < npos; pos += (int64_t)gridDim.x * PB) {
        float ym[8], d[8];
        unpack8(*reinterpret_cast<const uint4*>(ymax + pos * C + cg * 8), ym);
        unpack8(*reinterpret_cast<const uint4*>(da + pos * C + cg * 8), d);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const bool on = !relu || fmaf(ym[i], sc[i], shf[i]) > 0.f;
            const float dv = on ? d[i] : 0.f;
            acc1[i] += dv;
            acc2[i] = fmaf(dv, (ym[i] - mu[i]) * is[i], acc2[i]);
        }
    }
    float* r1 = sh_red;
    float* r2 = sh_red + PB * C;
#pragma unroll
    for (int i = 0; i < 8; ++i) { r1[pl * C + cg * 8 + i] = acc1[i]; r2[pl * C + cg * 8 + i] = acc2[i]; }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float a = 0.f, b = 0.f;
        for (int q = 0; q < PB; ++q) { a += r1[q * C + c]; b += r2[q * C + c]; }
        partials[((size_t)blockIdx.x * 2 + 0) * C + c] = a;
        partials[((size_t)blockIdx.x * 2 + 1) * C + c] = b;
    }
}

// apply: one thread = one 3x3x3 window (incl. the remainder windows at the far edges) x V channels.
//   dy = scale*dyhat + A*y + B   with  A = -scale*invstd*c2,  B = scale*(mean*invstd*c2 - c1)
template <int V>
__global__ void __launch_bounds__(256) bn_relu_pool3_bwd_apply_kernel(
    const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ ymax, const __nv_bfloat16* __restrict__ da,
    __nv_bfloat16* __restrict__ dy, BlockGeom g, const float* __restrict__ scale, const float* __restrict__ shift,
    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ c12, int relu) {
    using VT = typename VecT<V>::T;
    const int CG = g.C / V;
    const int Dw = (g.D + 2) / 3, Hw = (g.H + 2) / 3, Ww = (g.W + 2) / 3;
    const int64_t total = (int64_t)g.N * Dw * Hw * Ww * CG;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int cg = (int)(idx % CG);
    int64_t p = idx / CG;
    const int ww = (int)(p % Ww); p /= Ww;
    const int hw = (int)(p % Hw); p /= Hw;
    const int dw = (int)(p % Dw);
    const int n = (int)(p / Dw);
    float sc[V], shf[V], cA[V], cB[V], vb[V], dav[V];
    {
        float mu[V], is[V], k1[V], k2[V];
        ldv<V>(scale + cg * V, sc);
        ldv<V>(shift + cg * V, shf);
        ldv<V>(mean + cg * V, mu);
        ldv<V>(invstd + cg * V, is);
        ldv<V>(c12 + cg * V, k1);
        ldv<V>(c12 + g.C + cg * V, k2);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            cA[i] = -sc[i] * is[i] * k2[i];
            cB[i] = sc[i] * (mu[i] * is[i] * k2[i] - k1[i]);
        }
    }
    const bool has_out = dw < g.Dp && hw < g.Hp && ww < g.Wp;
#pragma unroll
    for (int i = 0; i < V; ++i) { vb[i] = INFINITY; dav[i] = 0.f; }
    if (has_out) {
        const int64_t po = ((((int64_t)n * g.Dp + dw) * g.Hp + hw) * g.Wp + ww) * g.C + cg * V;
        float ym[V];
        unpackv<V>(*reinterpret_cast<const VT*>(ymax + po), ym);
        unpackv<V>(*reinterpret_cast<const VT*>(da + po), dav);
#pragma unroll
        for (int i = 0; i < V; ++i) {
            vb[i] = fmaf(ym[i], sc[i], shf[i]);
            if (relu && !(vb[i] > 0.f)) { dav[i] = 0.f; vb[i] = INFINITY; }   // gated: nobody wins
            dav[i] *= sc[i];
        }
    }
    const int64_t sH = (int64_t)g.W * g.C, sD = (int64_t)g.H * g.W * g.C;
    const int64_t b0 = ((((int64_t)n * g.D + dw * 3) * g.H + hw * 3) * g.W + ww * 3) * g.C + cg * V;
    const bool interior = dw * 3 + 2 < g.D && hw * 3 + 2 < g.H && ww * 3 + 2 < g.W;
    if (interior) {
#pragma unroll
        for (int kd = 0; kd < 3; ++kd) {
            VT v[9];
#pragma unroll
            for (int t = 0; t < 9; ++t) v[t] = *reinterpret_cast<const VT*>(y + b0 + kd * sD + (t / 3) * sH + (t % 3) * g.C);
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                float f[V], o[V];
                unpackv<V>(v[t], f);
#pragma unroll
                for (int i = 0; i < V; ++i) {
                    const bool win = fmaf(f[i], sc[i], shf[i]) == vb[i];
                    o[i] = fmaf(cA[i], f[i], cB[i]) + (win ? dav[i] : 0.f);
                    if (win) vb[i] = INFINITY;       // first match only
                }
                *reinterpret_cast<VT*>(dy + b0 + kd * sD + (t / 3) * sH + (t % 3) * g.C) = packv<V>(o);
            }
        }
    } else {
        for (int kd = 0; kd < 3 && dw * 3 + kd < g.D; ++kd)
            for (int kh = 0; kh < 3 && hw * 3 + kh < g.H; ++kh)
                for (int kw = 0; kw < 3 && ww * 3 + kw < g.W; ++kw) {
                    const int64_t off = b0 + kd * sD + kh * sH + kw * g.C;
                    float f[V], o[V];
                    unpackv<V>(*reinterpret_cast<const VT*>(y + off), f);
#pragma unroll
                    for (int i = 0; i < V; ++i) {
                        const bool win = fmaf(f[i], sc[i], shf[i]) == vb[i];
                        o[i] = fmaf(cA[i], f[i], cB[i]) + (win ? dav[i] : 0.f);
                        if (win) vb[i] = INFINITY;
                    }
                    *reinterpret_cast<VT*>(dy + off) = packv<V>(o);
                }
    }
}

// partials [T][2][C] -> c12 [2][C] (c1 = sum(dyhat)/M, c2 = sum(dyhat*xhat)/M), dgamma, dbeta
// 16 channels x 64 row slices per CTA (as bn_finalize_kernel): with up to 592 partial rows a thread adds ~9 of them instead of 74,
// which is what this latency-bound kernel's 8 us were made of (420 launches per federated round)
__global__ void __launch_bounds__(kFinCh * kFinSlices) bn_bwd_finalize_kernel(const float* __restrict__ partials, int T, int C, double count,
                                                                              float* __restrict__ c12, float* __restrict__ dgamma,
                                                                              float* __restrict__ dbeta) {
    __shared__ double s1[kFinSlices][kFinCh], s2[kFinSlices][kFinCh];
    const int tx = threadIdx.x % kFinCh, ty = threadIdx.x / kFinCh;
    const int c = blockIdx.x * kFinCh + tx;
    double a = 0.0, b = 0.0;
    if (c < C) {
        int t = ty;
        for (; t + kFinSlices < T; t += 2 * kFinSlices) {          // 4 independent loads in flight
            const float v0 = partials[((size_t)t * 2 + 0) * C + c], v1 = partials[((size_t)t * 2 + 1) * C + c];
            const float v2 = partials[((size_t)(t + kFinSlices) * 2 + 0) * C + c], v3 = partials[((size_t)(t + kFinSlices) * 2 + 1) * C + c];
            a += (double)v0; b += (double)v1; a += (double)v2; b += (double)v3;
        }
        for (; t < T; t += kFinSlices) {
            a += (double)partials[((size_t)t * 2 + 0) * C + c];
            b += (double)partials[((size_t)t * 2 + 1) * C + c];
        }
    }
    s1[ty][tx] = a;
    s2[ty][tx] = b;
    __syncthreads();
    if (ty == 0 && c < C) {
        for (int i = 1; i < kFinSlices; ++i) { a += s1[i][tx]; b += s2[i][tx]; }
        c12[c] = (float)(a / count);
        c12[C + c] = (float)(b / count);
        if (dgamma) dgamma[c] = (float)b;
        if (dbeta) dbeta[c] = (float)a;
    }
}

}  // namespace nidt
using namespace nidt;

static inline BlockGeom make_geom(int N, int D, int H, int W, int C, int pk, int ps) {
    BlockGeom g{N, D, H, W, C, pk, ps, D, H, W};
    if (pk) { g.Dp = (D - pk) / ps + 1; g.Hp = (H - pk) / ps + 1; g.Wp = (W - pk) / ps + 1; }
    return g;
}

NIDT_API int nidt_bn_bwd_grid() { return 4 * kNumSMs; }

NIDT_API int nidt_bn_finalize(const float* partials, int T, int C, double count, const float* gamma, const float* beta,
                              float* running_mean, float* running_var, float* num_batches, float momentum, float eps,
                              float* scale, float* shift, float* mean, float* invstd, void* stream) {
    bn_finalize_kernel<<<(C + kFinCh - 1) / kFinCh, kFinCh * kFinSlices, 0, (cudaStream_t)stream>>>(partials, T, C, count, gamma, beta, running_mean, running_var,
                                                                        num_batches, momentum, eps, scale, shift, mean, invstd);
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_bn_eval_prepare(int C, const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                                  float eps, float* scale, float* shift, void* stream) {
    bn_eval_prepare_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(C, gamma, beta, running_mean, running_var, eps, scale, shift);
    NIDT_LAUNCH_CHECK();
    return 0;
}

static int bn_relu_pool_fwd_impl(const void* y, void* a, void* ymax, void* amax, int N, int D, int H, int W, int C, int pk, int ps,
                                 const float* scale, const float* shift, int relu, void* stream);
// ymax (bf16, pooled shape) is optional: when given with pool 3/3 the fast path runs and the backward can use it
NIDT_API int nidt_bn_relu_pool_fwd(const void* y, void* a, void* ymax, int N, int D, int H, int W, int C, int pk, int ps,
                                   const float* scale, const float* shift, int relu, void* stream) {
    return bn_relu_pool_fwd_impl(y, a, ymax, nullptr, N, D, H, W, C, pk, ps, scale, shift, relu, stream);
}
// the pool-3/3 fast path that also records, per pooled element, WHICH of the 27 window positions won (uint8 tap kd*9+kh*3+kw of the
// first maximum, 255 = gated by ReLU): the input of the fused BatchNorm-backward + weight-gradient kernel of the first layer
NIDT_API int nidt_bn_relu_pool3_fwd_amax(const void* y, void* a, void* ymax, void* amax, int N, int D, int H, int W, int C,
                                         const float* scale, const float* shift, int relu, void* stream) {
    NIDT_REQUIRE(ymax != nullptr && amax != nullptr, "nidt_bn_relu_pool3_fwd_amax: ymax / amax is null");
    return bn_relu_pool_fwd_impl(y, a, ymax, amax, N, D, H, W, C, 3, 3, scale, shift, relu, stream);
}
static int bn_relu_pool_fwd_impl(const void* y, void* a, void* ymax, void* amax, int N, int D, int H, int W, int C, int pk, int ps,
                                 const float* scale, const float* shift, int relu, void* stream) {
    NIDT_REQUIRE(C % 8 == 0, "nidt_bn_relu_pool_fwd: C=%d must be a multiple of 8", C);
    BlockGeom g = make_geom(N, D, H, W, C, pk, ps);
    const int64_t total = (int64_t)N * g.Dp * g.Hp * g.Wp * (C / 8);
    if (pk == 3 && ps == 3 && ymax != nullptr) {
        // 8 channels per thread measured faster than 4 for the forward (0.094 vs 0.110 ms on block 0 of AlexNet3D)
        if (amax != nullptr) {
            // with the winning taps tracked, 8 channels per thread need 109 registers (2 CTAs per SM): 4 channels keep the loads in flight
            const int64_t total4 = (int64_t)N * g.Dp * g.Hp * g.Wp * (C / 4);
            bn_relu_pool3_fwd_kernel<4, true><<<(unsigned)((total4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, (__nv_bfloat16*)a,
                                                                                                         (__nv_bfloat16*)ymax, g, scale, shift, relu,
                                                                                                         (unsigned char*)amax);
            NIDT_LAUNCH_CHECK();
            return 0;
        }
        static const int pool_v = [] { const char* e = getenv("NIDT_POOL_V"); return e ? atoi(e) : 8; }();
        if (pool_v == 4) {
            // 4 channels per thread at 80 registers (3 CTAs per SM) keeps more loads in flight than 8 channels at 106
            const int64_t total4 = (int64_t)N * g.Dp * g.Hp * g.Wp * (C / 4);
            bn_relu_pool3_fwd_kernel<4, false><<<(unsigned)((total4 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
                (const __nv_bfloat16*)y, (__nv_bfloat16*)a, (__nv_bfloat16*)ymax, g, scale, shift, relu, nullptr);
        } else
            bn_relu_pool3_fwd_kernel<8, false><<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, (__nv_bfloat16*)a,
                                                                                                    (__nv_bfloat16*)ymax, g, scale, shift, relu,
                                                                                                    (unsigned char*)amax);
        NIDT_LAUNCH_CHECK();
        return 0;
    }
    int64_t blocks = (total + 255) / 256;
    if (blocks > (int64_t)kNumSMs * 16) blocks = kNumSMs * 16;
    if (blocks < 1) blocks = 1;
    bn_relu_pool_fwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, (__nv_bfloat16*)a, g, scale, shift, relu);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// BatchNorm + ReLU + MaxPool3d(3,3) of a 64-channel conv output [N][D][H][W] that arrives reduced along W: yw bf16 [N][D][H][Wp][64]
// and aw uint16 [N][D][H][Wp][8] as nidt_conv_in_s2d_fprop_poolw writes them (Wp = (W - 3) / 3 + 1).  a: pooled activation
// [N][Dp][Hp][Wp][64]; ymax / amax (both or neither): raw conv output and tap (kd*9 + kh*3 + kw, 255 = gated by ReLU) of every
// window's winner -- the same three tensors nidt_bn_relu_pool3_fwd_amax produces from the full-resolution output.
NIDT_API int nidt_bn_relu_pool3w_fwd(const void* yw, const void* aw, void* a, void* ymax, void* amax, int N, int D, int H, int W, int C,
                                     const float* scale, const float* shift, int relu, void* stream) {
    NIDT_REQUIRE(C == 64 && D >= 3 && H >= 3 && W >= 3, "nidt_bn_relu_pool3w_fwd: C=%d (64) D=%d H=%d W=%d (>= 3)", C, D, H, W);
    NIDT_REQUIRE(yw != nullptr && aw != nullptr && a != nullptr && (ymax == nullptr) == (amax == nullptr), "nidt_bn_relu_pool3w_fwd: bad pointers");
    const int Dp = (D - 3) / 3 + 1, Hp = (H - 3) / 3 + 1, Wp = (W - 3) / 3 + 1;
    const int64_t total = (int64_t)N * Dp * Hp * Wp * 8;
    const unsigned grid = (unsigned)((total + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    if (ymax != nullptr)
        bn_relu_pool3w_fwd_kernel<true><<<grid, 256, 0, s>>>((const __nv_bfloat16*)yw, (const unsigned short*)aw, (__nv_bfloat16*)a, (__nv_bfloat16*)ymax,
                                                             (unsigned char*)amax, N, D, H, Dp, Hp, Wp, scale, shift, relu);
    else
        bn_relu_pool3w_fwd_kernel<false><<<grid, 256, 0, s>>>((const __nv_bfloat16*)yw, (const unsigned short*)aw, (__nv_bfloat16*)a, nullptr, nullptr, N,
                                                              D, H, Dp, Hp, Wp, scale, shift, relu);
    NIDT_LAUNCH_CHECK();
    return 0;
}

static int bn_relu_pool_bwd_impl(const void* y, const void* ymax, const void* da, void* dy, int N, int D, int H, int W, int C, int pk, int ps,
                                 const float* scale, const float* shift, const float* mean, const float* invstd,
                                 float* partials, float* c12, float* dgamma, float* dbeta, int relu, int phases, int* rows_out, void* stream);
constexpr int kBwdReduce = 1, kBwdFinalize = 2, kBwdApply = 4;
// partials: >= nidt_bn_bwd_grid()*2*C floats.  Sequence: reduce -> finalize -> apply.
NIDT_API int nidt_bn_relu_pool_bwd(const void* y, const void* ymax, const void* da, void* dy, int N, int D, int H, int W, int C, int pk, int ps,
                                   const float* scale, const float* shift, const float* mean, const float* invstd,
                                   float* partials, float* c12, float* dgamma, float* dbeta, int relu, void* stream) {
    return bn_relu_pool_bwd_impl(y, ymax, da, dy, N, D, H, W, C, pk, ps, scale, shift, mean, invstd, partials, c12, dgamma, dbeta, relu, kBwdReduce | kBwdFinalize | kBwdApply, nullptr, stream);
}
// The same three launches one at a time, for callers that put a cross-device sum between them (data-parallel training of one client:
// the per-channel partial sums are all-reduced over the ranks before the finalisation, ops/engine_dp.py).  phases: 1 = reduce (writes
// *rows_out rows of 2*C partial sums), 2 = finalise `rows` rows with `count` elements per channel, 4 = apply.
NIDT_API int nidt_bn_relu_pool_bwd_phases(const void* y, const void* ymax, const void* da, void* dy, int N, int D, int H, int W, int C, int pk,
                                          int ps, const float* scale, const float* shift, const float* mean, const float* invstd,
                                          float* partials, float* c12, float* dgamma, float* dbeta, int relu, int phases, int* rows_out,
                                          void* stream) {
    NIDT_REQUIRE(phases == kBwdReduce || phases == kBwdApply, "nidt_bn_relu_pool_bwd_phases: phases=%d (1 = reduce, 4 = apply)", phases);
    return bn_relu_pool_bwd_impl(y, ymax, da, dy, N, D, H, W, C, pk, ps, scale, shift, mean, invstd, partials, c12, dgamma, dbeta, relu, phases,
                                 rows_out, stream);
}
NIDT_API int nidt_bn_bwd_finalize(const float* partials, int rows, int C, double count, float* c12, float* dgamma, float* dbeta, void* stream) {
    NIDT_REQUIRE(rows >= 1 && C >= 1 && count > 0 && c12 != nullptr, "nidt_bn_bwd_finalize: rows=%d C=%d", rows, C);
    bn_bwd_finalize_kernel<<<(C + kFinCh - 1) / kFinCh, kFinCh * kFinSlices, 0, (cudaStream_t)stream>>>(partials, rows, C, count, c12, dgamma, dbeta);
    NIDT_LAUNCH_CHECK();
    return 0;
}
// pool 3/3 with ymax: only the reductions (c12 = the two BatchNorm-backward means, dgamma, dbeta) -- dY itself is then formed on the
// fly by the consumer (nidt_conv_in_s2d_wgrad_fused) instead of being written to and read back from HBM
NIDT_API int nidt_bn_relu_pool3_bwd_stats(const void* ymax, const void* da, int N, int D, int H, int W, int C, const float* scale,
                                          const float* shift, const float* mean, const float* invstd, float* partials, float* c12,
                                          float* dgamma, float* dbeta, int relu, void* stream) {
    NIDT_REQUIRE(ymax != nullptr, "nidt_bn_relu_pool3_bwd_stats: ymax is null");
    return bn_relu_pool_bwd_impl(nullptr, ymax, da, nullptr, N, D, H, W, C, 3, 3, scale, shift, mean, invstd, partials, c12, dgamma, dbeta, relu, kBwdReduce | kBwdFinalize, nullptr, stream);
}
static int bn_relu_pool_bwd_impl(const void* y, const void* ymax, const void* da, void* dy, int N, int D, int H, int W, int C, int pk, int ps,
                                 const float* scale, const float* shift, const float* mean, const float* invstd,
                                 float* partials, float* c12, float* dgamma, float* dbeta, int relu, int phases, int* rows_out, void* stream) {
    NIDT_REQUIRE(C % 8 == 0 && C <= 2048, "nidt_bn_relu_pool_bwd: unsupported C=%d", C);
    BlockGeom g = make_geom(N, D, H, W, C, pk, ps);
    const int CG = C / 8;
    const int PB = 256 / CG > 0 ? 256 / CG : 1;
    const int threads = PB * CG;
    NIDT_REQUIRE(threads <= 1024, "nidt_bn_relu_pool_bwd: C too large");
    const int grid = nidt_bn_bwd_grid();
    const size_t smem = (size_t)2 * PB * C * sizeof(float);
    cudaStream_t s = (cudaStream_t)stream;
    const double count = (double)N * D * H * W;
    if (pk == 3 && ps == 3 && ymax != nullptr) {
        const int64_t npos = (int64_t)N * g.Dp * g.Hp * g.Wp;
        int rgrid = (int)((npos + PB - 1) / PB);
        if (rgrid > grid) rgrid = grid;
        if (phases & kBwdReduce) {
            bn_pool3_bwd_reduce_kernel<<<rgrid, threads, smem, s>>>((const __nv_bfloat16*)ymax, (const __nv_bfloat16*)da, npos, C, scale, shift, mean,
                                                                   invstd, partials, relu);
            NIDT_LAUNCH_CHECK();
            if (rows_out) *rows_out = rgrid;
        }
        if (phases & kBwdFinalize) {
            bn_bwd_finalize_kernel<<<(C + kFinCh - 1) / kFinCh, kFinCh * kFinSlices, 0, s>>>(partials, rgrid, C, count, c12, dgamma, dbeta);
            NIDT_LAUNCH_CHECK();
        }
        if (!(phases & kBwdApply)) return 0;
        static const int bwd_v = [] { const char* e = getenv("NIDT_BWD_V"); return e ? atoi(e) : 4; }();
        if (bwd_v == 8) {
            const int64_t total8 = (int64_t)N * ((D + 2) / 3) * ((H + 2) / 3) * ((W + 2) / 3) * (C / 8);
            bn_relu_pool3_bwd_apply_kernel<8><<<(unsigned)((total8 + 255) / 256), 256, 0, s>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)ymax,
                                                                                            (const __nv_bfloat16*)da, (__nv_bfloat16*)dy, g, scale, shift,
                                                                                            mean, invstd, c12, relu);
            NIDT_LAUNCH_CHECK();
            return 0;
        }
        const int64_t total = (int64_t)N * ((D + 2) / 3) * ((H + 2) / 3) * ((W + 2) / 3) * (C / 4);
        bn_relu_pool3_bwd_apply_kernel<4><<<(unsigned)((total + 255) / 256), 256, 0, s>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)ymax,
                                                                                       (const __nv_bfloat16*)da, (__nv_bfloat16*)dy, g, scale, shift,
                                                                                       mean, invstd, c12, relu);
        NIDT_LAUNCH_CHECK();
        return 0;
    }
    if (phases & kBwdReduce) {
        bn_relu_pool_bwd_kernel<false><<<grid, threads, smem, s>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)da, nullptr, g, scale, shift,
                                                                   mean, invstd, nullptr, partials, relu);
        NIDT_LAUNCH_CHECK();
        if (rows_out) *rows_out = grid;
    }
    if (phases & kBwdFinalize) {
        bn_bwd_finalize_kernel<<<(C + kFinCh - 1) / kFinCh, kFinCh * kFinSlices, 0, s>>>(partials, grid, C, count, c12, dgamma, dbeta);
        NIDT_LAUNCH_CHECK();
    }
    if (!(phases & kBwdApply)) return 0;
    bn_relu_pool_bwd_kernel<true><<<grid, threads, 0, s>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)da, (__nv_bfloat16*)dy, g, scale, shift,
                                                           mean, invstd, c12, nullptr, relu);
    NIDT_LAUNCH_CHECK();
    return 0;
}
