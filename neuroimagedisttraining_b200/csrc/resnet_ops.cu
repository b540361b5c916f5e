// Glue kernels of the native 3-D ResNet path (BASELINE config 4; reference model: ResNet_l3 with BasicBlock,
// fedml_api/model/cv/salient_models.py:13-139).  The convolutions and BatchNorms are the tcgen05 / bn_pool kernels the
// AlexNet path uses; this file adds what a residual network needs on top, all on channels-last bf16 activations:
//   residual add + ReLU (forward) and its gate (backward), gradient accumulation at the block input,
//   MaxPool3d with overlapping windows and padding (k3 s2 p1) forward (+ tap index) / backward (gather form, no atomics),
//   zero-insertion upsampling (turns the data gradient of a stride-s conv into a stride-1 conv the dgrad kernel runs),
//   AvgPool3d forward / backward, and the Linear(F,Hd) -> Linear(Hd,1) -> BCEWithLogits head with its full backward.
// In the reference these are ATen / cuDNN / cuBLAS library calls driven by autograd.
#include "common.cuh"

namespace nidt {
namespace {

__device__ __forceinline__ void unpack8(uint4 v, float* f) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 t = __bfloat1622float2(h[i]);
        f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
    uint4 v;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return v;
}

inline int grid_for(int64_t work, int threads) {
    int64_t b = (work + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * 16;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

// ---------------------------------------------------------------------------- elementwise (8 bf16 per thread)
// mode 0: out = relu(a + b)      mode 1: out = (b > 0) ? a : 0  (a = dout, b = forward output)      mode 2: out = a + b
template <int kMode>
__global__ void __launch_bounds__(256) ew_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ out, int64_t n8) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        float x[8], y[8];
        unpack8(a[i], x);
        unpack8(b[i], y);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (kMode == 0) x[j] = fmaxf(x[j] + y[j], 0.f);
            else if (kMode == 1) x[j] = y[j] > 0.f ? x[j] : 0.f;
            else x[j] = x[j] + y[j];
        }
        out[i] = pack8(x);
    }
}

struct PoolGeom {
    int N, D, H, W, C;      // input
    int Do, Ho, Wo;         // output
    int k, s, p;
};

// ---------------------------------------------------------------------------- MaxPool3d (general k / stride / padding)
// one thread = one output position x 8 channels; idx = tap (kd*k+kh)*k+kw of the first maximum in scan order (PyTorch's rule)
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                          uint8_t* __restrict__ idx, PoolGeom g) {
    const int CG = g.C >> 3;
    const int64_t total = (int64_t)g.N * g.Do * g.Ho * g.Wo * CG;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int cg = (int)(t % CG);
        int64_t q = t / CG;
        const int wo = (int)(q % g.Wo); q /= g.Wo;
        const int ho = (int)(q % g.Ho); q /= g.Ho;
        const int od = (int)(q % g.Do);
        const int n = (int)(q / g.Do);
        float best[8];
        int arg[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
        for (int kd = 0; kd < g.k; ++kd) {
            const int d = od * g.s - g.p + kd;
            if (d < 0 || d >= g.D) continue;
            for (int kh = 0; kh < g.k; ++kh) {
                const int h = ho * g.s - g.p + kh;
                if (h < 0 || h >= g.H) continue;
                for (int kw = 0; kw < g.k; ++kw) {
                    const int w = wo * g.s - g.p + kw;
                    if (w < 0 || w >= g.W) continue;
                    float f[8];
                    unpack8(*reinterpret_cast<const uint4*>(x + ((((int64_t)n * g.D + d) * g.H + h) * g.W + w) * g.C + cg * 8), f);
                    const int tap = (kd * g.k + kh) * g.k + kw;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (f[j] > best[j]) { best[j] = f[j]; arg[j] = tap; }
                }
            }
        }
        *reinterpret_cast<uint4*>(out + t * 8) = pack8(best);
        if (idx != nullptr) {
            uint2 packed;
            packed.x = (uint32_t)arg[0] | ((uint32_t)arg[1] << 8) | ((uint32_t)arg[2] << 16) | ((uint32_t)arg[3] << 24);
            packed.y = (uint32_t)arg[4] | ((uint32_t)arg[5] << 8) | ((uint32_t)arg[6] << 16) | ((uint32_t)arg[7] << 24);
            *reinterpret_cast<uint2*>(idx + t * 8) = packed;
        }
    }
}

// one thread = one INPUT position x 8 channels: sums dout of every window that covers it and elected it (no atomics)
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dout, const uint8_t* __restrict__ idx,
                                                          __nv_bfloat16* __restrict__ dx, PoolGeom g) {
    const int CG = g.C >> 3;
    const int64_t total = (int64_t)g.N * g.D * g.H * g.W * CG;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int cg = (int)(t % CG);
        int64_t q = t / CG;
        const int w = (int)(q % g.W); q /= g.W;
        const int h = (int)(q % g.H); q /= g.H;
        const int d = (int)(q % g.D);
        const int n = (int)(q / g.D);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        // windows o with o*s - p <= pos <= o*s - p + k - 1
        const int od_lo = max(0, (d + g.p - g.k + g.s) / g.s), od_hi = min(g.Do - 1, (d + g.p) / g.s);
        const int oh_lo = max(0, (h + g.p - g.k + g.s) / g.s), oh_hi = min(g.Ho - 1, (h + g.p) / g.s);
        const int ow_lo = max(0, (w + g.p - g.k + g.s) / g.s), ow_hi = min(g.Wo - 1, (w + g.p) / g.s);
        for (int od = od_lo; od <= od_hi; ++od)
            for (int oh = oh_lo; oh <= oh_hi; ++oh)
                for (int ow = ow_lo; ow <= ow_hi; ++ow) {
                    const int tap = ((d + g.p - od * g.s) * g.k + (h + g.p - oh * g.s)) * g.k + (w + g.p - ow * g.s);
                    const int64_t o = ((((int64_t)n * g.Do + od) * g.Ho + oh) * g.Wo + ow) * g.C + cg * 8;
                    const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
                    float f[8];
                    unpack8(*reinterpret_cast<const uint4*>(dout + o), f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint32_t a = ((j < 4 ? pk.x : pk.y) >> (8 * (j & 3))) & 0xffu;
                        if ((int)a == tap) acc[j] += f[j];
                    }
                }
        *reinterpret_cast<uint4*>(dx + t * 8) = pack8(acc);
    }
}

// ---------------------------------------------------------------------------- zero-insertion upsampling
// u[n, s*d, s*h, s*w, :] = dy[n, d, h, w, :], 0 elsewhere; u has extents (Du, Hu, Wu) >= s*(D-1)+1
__global__ void __launch_bounds__(256) upsample_zero_kernel(const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ u, int N, int D, int H,
                                                            int W, int C, int Du, int Hu, int Wu, int s) {
    const int CG = C >> 3;
    const int64_t total = (int64_t)N * Du * Hu * Wu * CG;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int cg = (int)(t % CG);
        int64_t q = t / CG;
        const int w = (int)(q % Wu); q /= Wu;
        const int h = (int)(q % Hu); q /= Hu;
        const int d = (int)(q % Du);
        const int n = (int)(q / Du);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (d % s == 0 && h % s == 0 && w % s == 0 && d / s < D && h / s < H && w / s < W)
            v = *reinterpret_cast<const uint4*>(dy + ((((int64_t)n * D + d / s) * H + h / s) * W + w / s) * C + cg * 8);
        *reinterpret_cast<uint4*>(u + t * 8) = v;
    }
}

// ---------------------------------------------------------------------------- AvgPool3d(k, stride k)
// feat[n][c*S + s] (fp32, the NCDHW flatten order the reference's fc weights expect), S = Do*Ho*Wo
__global__ void __launch_bounds__(256) avgpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ feat, PoolGeom g) {
    const int S = g.Do * g.Ho * g.Wo;
    const int64_t total = (int64_t)g.N * S * g.C;
    const float inv = 1.f / (float)(g.k * g.k * g.k);
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(t % g.C);
        int64_t q = t / g.C;
        const int sidx = (int)(q % S);
        const int n = (int)(q / S);
        const int wo = sidx % g.Wo, ho = (sidx / g.Wo) % g.Ho, od = sidx / (g.Wo * g.Ho);
        float acc = 0.f;
        for (int kd = 0; kd < g.k; ++kd)
            for (int kh = 0; kh < g.k; ++kh)
                for (int kw = 0; kw < g.k; ++kw)
                    acc += __bfloat162float(x[((((int64_t)n * g.D + od * g.k + kd) * g.H + ho * g.k + kh) * g.W + wo * g.k + kw) * g.C + c]);
        feat[(int64_t)n * S * g.C + (int64_t)c * S + sidx] = acc * inv;
    }
}

__global__ void __launch_bounds__(256) avgpool_bwd_kernel(const float* __restrict__ dfeat, __nv_bfloat16* __restrict__ dx, PoolGeom g) {
    const int S = g.Do * g.Ho * g.Wo;
    const int64_t total = (int64_t)g.N * g.D * g.H * g.W * g.C;
    const float inv = 1.f / (float)(g.k * g.k * g.k);
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(t % g.C);
        int64_t q = t / g.C;
        const int w = (int)(q % g.W); q /= g.W;
        const int h = (int)(q % g.H); q /= g.H;
        const int d = (int)(q % g.D);
        const int n = (int)(q / g.D);
        const int od = d / g.k, ho = h / g.k, wo = w / g.k;
        float v = 0.f;
        if (od < g.Do && ho < g.Ho && wo < g.Wo) v = dfeat[(int64_t)n * S * g.C + (int64_t)c * S + (od * g.Ho + ho) * g.Wo + wo] * inv;
        dx[t] = __float2bfloat16(v);
    }
}

// ---------------------------------------------------------------------------- Linear(F,Hd) -> Linear(Hd,1) -> BCE head
constexpr int kHeadMaxB = 32;

// h[n][j] = b1[j] + sum_k W1[j][k] f[n][k]; one warp per hidden unit j
__global__ void __launch_bounds__(256) lin_fwd_kernel(const float* __restrict__ f, const float* __restrict__ W1, const float* __restrict__ b1,
                                                      float* __restrict__ h, int B, int F, int Hd) {
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (j >= Hd) return;
    float acc[kHeadMaxB];
#pragma unroll
    for (int n = 0; n < kHeadMaxB; ++n) acc[n] = 0.f;
    for (int k = lane; k < F; k += 32) {
        const float w = W1[(int64_t)j * F + k];
#pragma unroll
        for (int n = 0; n < kHeadMaxB; ++n)
            if (n < B) acc[n] = fmaf(w, f[(int64_t)n * F + k], acc[n]);
    }
#pragma unroll
    for (int n = 0; n < kHeadMaxB; ++n)
        if (n < B) {
            const float s = warp_sum(acc[n]);
            if (lane == 0) h[(int64_t)n * Hd + j] = s + b1[j];
        }
}

struct LinHeadArgs {
    const float *y, *W2, *b2, *h;
    float *gW2, *gb2, *gb1, *dh;
    float* loss_acc;
    double* eval_counters;
    float* logits_out;
    int B, Hd, train, double_sigmoid;
};

// one CTA: logits, loss / eval counters, dz, gW2, gb2, dh = dz (x) W2, gb1
__global__ void __launch_bounds__(512) lin_mid_kernel(LinHeadArgs a) {
    __shared__ float z[kHeadMaxB], dz[kHeadMaxB];
    const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    for (int n = wid; n < a.B; n += blockDim.x >> 5) {
        float acc = 0.f;
        for (int j = lane; j < a.Hd; j += 32) acc = fmaf(a.W2[j], a.h[(int64_t)n * a.Hd + j], acc);
        acc = warp_sum(acc);
        if (lane == 0) {
            acc += a.b2[0];
            z[n] = acc;
            if (a.logits_out) a.logits_out[n] = acc;
            dz[n] = (1.f / (1.f + __expf(-acc)) - a.y[n]) / (float)a.B;
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (a.loss_acc != nullptr) {
            float l = 0.f;
            for (int n = 0; n < a.B; ++n) {
                const float x = z[n], t = a.y[n];
                l += fmaxf(x, 0.f) - x * t + log1pf(__expf(-fabsf(x)));
            }
            *a.loss_acc += l / (float)a.B;
        }
        if (a.eval_counters != nullptr) {
            double correct = 0.0, loss = 0.0;
            for (int n = 0; n < a.B; ++n) {
                const float t = a.y[n];
                const float pr = 1.f / (1.f + expf(-z[n]));
                const float x = a.double_sigmoid ? pr : z[n];
                loss += (double)(fmaxf(x, 0.f) - x * t + log1pf(expf(-fabsf(x))));
                correct += ((pr >= 0.5f ? 1.f : 0.f) == t) ? 1.0 : 0.0;
            }
            atomicAdd(&a.eval_counters[0], correct);
            atomicAdd(&a.eval_counters[1], loss);
            atomicAdd(&a.eval_counters[2], (double)a.B);
        }
    }
    if (!a.train) return;
    for (int j = tid; j < a.Hd; j += blockDim.x) {
        float gw = 0.f, gb = 0.f;
        const float w2 = a.W2[j];
        for (int n = 0; n < a.B; ++n) {
            gw = fmaf(dz[n], a.h[(int64_t)n * a.Hd + j], gw);
            const float d = dz[n] * w2;
            a.dh[(int64_t)n * a.Hd + j] = d;
            gb += d;
        }
        a.gW2[j] = gw;
        a.gb1[j] = gb;
    }
    if (tid == 0) {
        float g = 0.f;
        for (int n = 0; n < a.B; ++n) g += dz[n];
        a.gb2[0] = g;
    }
}

// gW1[j][k] = sum_n dh[n][j] f[n][k]
__global__ void __launch_bounds__(256) lin_bwd_w_kernel(const float* __restrict__ dh, const float* __restrict__ f, float* __restrict__ gW1, int B, int F, int Hd) {
    const int64_t total = (int64_t)Hd * F;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int j = (int)(t / F), k = (int)(t % F);
        float g = 0.f;
        for (int n = 0; n < B; ++n) g = fmaf(dh[(int64_t)n * Hd + j], f[(int64_t)n * F + k], g);
        gW1[t] = g;
    }
}
// df[n][k] = sum_j dh[n][j] W1[j][k]: a block owns 32 columns k; 8 warps split j, every W1 element is read once
__global__ void __launch_bounds__(256) lin_bwd_x_kernel(const float* __restrict__ dh, const float* __restrict__ W1, float* __restrict__ df, int B, int F, int Hd) {
    __shared__ float red[8][kHeadMaxB][33];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int k = blockIdx.x * 32 + lane;
    float acc[kHeadMaxB];
#pragma unroll
    for (int n = 0; n < kHeadMaxB; ++n) acc[n] = 0.f;
    if (k < F)
        for (int j = wid; j < Hd; j += 8) {
            const float w = W1[(int64_t)j * F + k];
#pragma unroll
            for (int n = 0; n < kHeadMaxB; ++n)
                if (n < B) acc[n] = fmaf(dh[(int64_t)n * Hd + j], w, acc[n]);
        }
#pragma unroll
    for (int n = 0; n < kHeadMaxB; ++n)
        if (n < B) red[wid][n][lane] = acc[n];
    __syncthreads();
    for (int i = threadIdx.x; i < B * 32; i += blockDim.x) {
        const int n = i >> 5, l = i & 31;
        if (blockIdx.x * 32 + l < F) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += red[w][n][l];
            df[(int64_t)n * F + blockIdx.x * 32 + l] = s;
        }
    }
}

}  // namespace
}  // namespace nidt
using namespace nidt;

// mode 0: out = relu(a + b); 1: out = b > 0 ? a : 0; 2: out = a + b.  n bf16 elements, n % 8 == 0; out may alias a.
NIDT_API int nidt_ew_bf16(const void* a, const void* b, void* out, int64_t n, int mode, void* stream) {
    NIDT_REQUIRE(n % 8 == 0 && mode >= 0 && mode <= 2, "nidt_ew_bf16: n=%lld must be a multiple of 8, mode in 0..2", (long long)n);
    const int64_t n8 = n / 8;
    const int grid = grid_for(n8, 256);
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == 0) ew_kernel<0><<<grid, 256, 0, s>>>((const uint4*)a, (const uint4*)b, (uint4*)out, n8);
    else if (mode == 1) ew_kernel<1><<<grid, 256, 0, s>>>((const uint4*)a, (const uint4*)b, (uint4*)out, n8);
    else ew_kernel<2><<<grid, 256, 0, s>>>((const uint4*)a, (const uint4*)b, (uint4*)out, n8);
    NIDT_LAUNCH_CHECK();
    return 0;
}

static int pool_geom(PoolGeom& g, int N, int D, int H, int W, int C, int k, int s, int p) {
    NIDT_REQUIRE(C % 8 == 0 && k >= 1 && k <= 6 && s >= 1 && p >= 0 && 2 * p <= k, "pool: unsupported C=%d k=%d s=%d p=%d", C, k, s, p);
    g.N = N; g.D = D; g.H = H; g.W = W; g.C = C; g.k = k; g.s = s; g.p = p;
    g.Do = (D + 2 * p - k) / s + 1; g.Ho = (H + 2 * p - k) / s + 1; g.Wo = (W + 2 * p - k) / s + 1;
    NIDT_REQUIRE(g.Do >= 1 && g.Ho >= 1 && g.Wo >= 1, "pool: input %dx%dx%d too small", D, H, W);
    return 0;
}

// MaxPool3d(k, s, p), floor mode.  idx: uint8 per output element (may be null for inference).  ref: salient_models.py:96 (ATen).
NIDT_API int nidt_maxpool3d_fwd(const void* x, void* out, void* idx, int N, int D, int H, int W, int C, int k, int s, int p, void* stream) {
    PoolGeom g;
    if (int rc = pool_geom(g, N, D, H, W, C, k, s, p)) return rc;
    const int64_t work = (int64_t)N * g.Do * g.Ho * g.Wo * (C / 8);
    maxpool_fwd_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, (uint8_t*)idx, g);
    NIDT_LAUNCH_CHECK();
    return 0;
}
NIDT_API int nidt_maxpool3d_bwd(const void* dout, const void* idx, void* dx, int N, int D, int H, int W, int C, int k, int s, int p, void* stream) {
    PoolGeom g;
    if (int rc = pool_geom(g, N, D, H, W, C, k, s, p)) return rc;
    const int64_t work = (int64_t)N * D * H * W * (C / 8);
    maxpool_bwd_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dout, (const uint8_t*)idx, (__nv_bfloat16*)dx, g);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// u (extents Du,Hu,Wu) = dy with s-1 zeros inserted between neighbours (and zero tails).
NIDT_API int nidt_upsample_zero(const void* dy, void* u, int N, int D, int H, int W, int C, int Du, int Hu, int Wu, int s, void* stream) {
    NIDT_REQUIRE(C % 8 == 0 && s >= 1 && Du >= s * (D - 1) + 1 && Hu >= s * (H - 1) + 1 && Wu >= s * (W - 1) + 1, "nidt_upsample_zero: bad extents");
    const int64_t work = (int64_t)N * Du * Hu * Wu * (C / 8);
    upsample_zero_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)dy, (__nv_bfloat16*)u, N, D, H, W, C, Du, Hu, Wu, s);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// AvgPool3d(k) (stride k, floor): bf16 NDHWC -> fp32 features in NCDHW flatten order, and back.  ref: salient_models.py:100,113.
NIDT_API int nidt_avgpool3d_fwd(const void* x, float* feat, int N, int D, int H, int W, int C, int k, void* stream) {
    PoolGeom g;
    if (int rc = pool_geom(g, N, D, H, W, C, k, k, 0)) return rc;
    const int64_t work = (int64_t)N * g.Do * g.Ho * g.Wo * C;
    avgpool_fwd_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, feat, g);
    NIDT_LAUNCH_CHECK();
    return 0;
}
NIDT_API int nidt_avgpool3d_bwd(const float* dfeat, void* dx, int N, int D, int H, int W, int C, int k, void* stream) {
    PoolGeom g;
    if (int rc = pool_geom(g, N, D, H, W, C, k, k, 0)) return rc;
    const int64_t work = (int64_t)N * D * H * W * C;
    avgpool_bwd_kernel<<<grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(dfeat, (__nv_bfloat16*)dx, g);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// feat [B][F] fp32 -> x1 = W1 feat + b1 (hidden, returned in `hidden`) -> logit = W2 x1 + b2 -> BCEWithLogits (mean), and the
// complete backward (gW1, gb1, gW2, gb2 into the gradient slab, dfeat [B][F]).  Eval: counters as nidt_head_fwd_bwd.
// hidden / dhidden: caller scratch of B*Hd floats each.  ref: salient_models.py:99-101,114-116; my_model_trainer.py:207-222,239-274.
NIDT_API int nidt_linear_head_fwd_bwd(const float* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                                      float* gW1, float* gb1, float* gW2, float* gb2, float* dfeat, float* hidden, float* dhidden,
                                      float* loss_acc, double* eval_counters, float* logits_out, int B, int F, int Hd, int train,
                                      int double_sigmoid, void* stream) {
    NIDT_REQUIRE(B >= 1 && B <= kHeadMaxB, "nidt_linear_head_fwd_bwd: batch %d > %d", B, kHeadMaxB);
    NIDT_REQUIRE(hidden != nullptr && (!train || dhidden != nullptr), "nidt_linear_head_fwd_bwd: scratch missing");
    cudaStream_t s = (cudaStream_t)stream;
    lin_fwd_kernel<<<(Hd + 7) / 8, 256, 0, s>>>(feat, W1, b1, hidden, B, F, Hd);
    NIDT_LAUNCH_CHECK();
    const bool bwd = train && gW1 != nullptr;
    LinHeadArgs a{y, W2, b2, hidden, gW2, gb2, gb1, dhidden, loss_acc, eval_counters, logits_out, B, Hd, bwd ? 1 : 0, double_sigmoid};
    lin_mid_kernel<<<1, 512, 0, s>>>(a);
    NIDT_LAUNCH_CHECK();
    if (bwd) {
        lin_bwd_w_kernel<<<grid_for((int64_t)Hd * F, 256), 256, 0, s>>>(dhidden, feat, gW1, B, F, Hd);
        NIDT_LAUNCH_CHECK();
        if (dfeat != nullptr) {
            lin_bwd_x_kernel<<<(F + 31) / 32, 256, 0, s>>>(dhidden, W1, dfeat, B, F, Hd);
            NIDT_LAUNCH_CHECK();
        }
    }
    return 0;
}
