// Classifier head of the 3-D AlexNets fused into one launch:
//   Dropout(p) -> Linear(F,Hd) -> ReLU -> Dropout(p) -> Linear(Hd,1) -> BCEWithLogits(mean)
// forward, loss and the complete backward (dW1, db1, dW2, db2 written straight into the fp32
// gradient slab, d(features) written as bf16 in the channels-last layout of the last pooled
// activation).  Replaces K6+K7 of SURVEY §2.5 (2 dropout, 2 cuBLAS GEMMs, ReLU, loss and their
// backward = ~20 library launches; reference: salient_models.py:171-176, my_model_trainer.py:207-222).
// Also the eval head (K16): sigmoid, threshold 0.5, (double-sigmoid) BCE, device-side counters.
//
// Feature order: the reference flattens NCDHW (feature = c*S + s); activations here are NDHWC
// (memory = s*C + c).  The kernel maps between the two so state_dict weights stay interchangeable.
#include "common.cuh"
#include "../../include/nidt_b200.h"

namespace nidt {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
// counter-based uniform in [0,1): (seed, stream id, element index)
__device__ __forceinline__ float uniform01(uint64_t seed, uint32_t sid, uint32_t idx) {
    uint32_t h = mix32((uint32_t)seed ^ mix32(idx * 0x9E3779B9U + sid));
    h = mix32(h ^ (uint32_t)(seed >> 32) ^ (sid * 0x85ebca6bU));
    return (float)(h >> 8) * (1.0f / 16777216.0f);
}

struct HeadArgs {
    const __nv_bfloat16* feat;   // [B][S][C]
    const float* y;              // [B]
    const float *W1, *b1, *W2, *b2;
    float *gW1, *gb1, *gW2, *gb2;   // may be null (eval)
    __nv_bfloat16* dfeat;        // [B][S][C], may be null
    float* loss_acc;             // += mean loss (train) ; may be null
    double* eval_counters;       // [3] correct, loss_sum, total (eval); may be null
    float* logits_out;           // [B], may be null
    int B, S, C, Hd;
    float p_drop;
    uint64_t seed;
    int train;                   // dropout on / backward
    int double_sigmoid;          // eval loss = BCEWithLogits(sigmoid(z)) as the reference does
    const nidt_dyn_t* dyn;       // when set: seed, loss_acc (train) / eval_counters (eval) come from device memory (graph replays)
};

// one CTA, 1024 threads; dynamic smem: f[B][F] (post-dropout, fp32) | h[B][Hd] | dh[B][Hd] | z[B] | dz[B] | W1[Hd][F+1]
// (W1 is staged once with coalesced loads; every later pass reads it from shared memory)
__global__ void __launch_bounds__(1024) head_kernel(const __grid_constant__ HeadArgs args) {
    extern __shared__ float sm[];
    HeadArgs a = args;
    if (a.dyn != nullptr) {
        a.seed = a.dyn->seed;
        if (a.train) a.loss_acc = a.dyn->loss_acc;
        else a.eval_counters = a.dyn->counters;
    }
    const int F = a.S * a.C;
    float* f = sm;
    float* h = f + (size_t)a.B * F;
    float* dh = h + (size_t)a.B * a.Hd;
    float* z = dh + (size_t)a.B * a.Hd;
    float* dz = z + a.B;
    float* sW1 = dz + a.B;
    const int FP = F + 1;
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < a.Hd * F; i += nt) sW1[(i / F) * FP + (i % F)] = a.W1[i];
    const float keep_scale = (a.train && a.p_drop > 0.f) ? 1.f / (1.f - a.p_drop) : 1.f;

    // features: memory (n, s, c) -> logical index c*S + s ; dropout 1
    for (int i = tid; i < a.B * F; i += nt) {
        const int n = i / F, r = i % F;
        const int s = r / a.C, c = r % a.C;
        float v = __bfloat162float(a.feat[i]);
        const int lf = c * a.S + s;
        if (a.train && a.p_drop > 0.f) v = (uniform01(a.seed, 1u, (uint32_t)(n * F + lf)) >= a.p_drop) ? v * keep_scale : 0.f;
        f[n * F + lf] = v;
    }
    __syncthreads();
    // fc1 + relu + dropout 2:  h[n][j]
    for (int i = tid; i < a.B * a.Hd; i += nt) {
        const int n = i / a.Hd, j = i % a.Hd;
        const float* w = sW1 + (size_t)j * FP;
        const float* x = f + (size_t)n * F;
        float acc = a.b1[j];
        for (int k = 0; k < F; ++k) acc = fmaf(w[k], x[k], acc);
        float v = fmaxf(acc, 0.f);
        float m = (acc > 0.f) ? 1.f : 0.f;
        if (a.train && a.p_drop > 0.f) {
            const float keep = (uniform01(a.seed, 2u, (uint32_t)i) >= a.p_drop) ? keep_scale : 0.f;
            v *= keep; m *= keep;
        }
        h[i] = v;
        dh[i] = m;     // d h / d pre-activation (relu' * dropout scale), reused in backward
    }
    __syncthreads();
    // fc2 -> logits
    if (tid < a.B) {
        float acc = a.b2[0];
        for (int j = 0; j < a.Hd; ++j) acc = fmaf(a.W2[j], h[tid * a.Hd + j], acc);
        z[tid] = acc;
        if (a.logits_out) a.logits_out[tid] = acc;
        const float t = a.y[tid];
        const float sg = 1.f / (1.f + __expf(-acc));
        dz[tid] = (sg - t) / (float)a.B;
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
    if (!a.train || a.gW1 == nullptr) return;

    // ---- backward
    // gW2[j] = sum_n dz[n]*h[n][j] ; gb2 = sum dz ; dpre[n][j] = dz[n]*W2[j]*dh[n][j]
    for (int j = tid; j < a.Hd; j += nt) {
        float g = 0.f;
        for (int n = 0; n < a.B; ++n) g = fmaf(dz[n], h[n * a.Hd + j], g);
        a.gW2[j] = g;
    }
    if (tid == 0) {
        float g = 0.f;
        for (int n = 0; n < a.B; ++n) g += dz[n];
        a.gb2[0] = g;
    }
    __syncthreads();
    for (int i = tid; i < a.B * a.Hd; i += nt) {
        const int n = i / a.Hd, j = i % a.Hd;
        dh[i] = dz[n] * a.W2[j] * dh[i];      // now d loss / d pre-activation of fc1
    }
    __syncthreads();
    for (int j = tid; j < a.Hd; j += nt) {
        float g = 0.f;
        for (int n = 0; n < a.B; ++n) g += dh[n * a.Hd + j];
        a.gb1[j] = g;
    }
    // gW1[j][k] = sum_n dpre[n][j] * f[n][k]
    for (int i = tid; i < a.Hd * F; i += nt) {
        const int j = i / F, k = i % F;
        float g = 0.f;
        for (int n = 0; n < a.B; ++n) g = fmaf(dh[n * a.Hd + j], f[n * F + k], g);
        a.gW1[i] = g;
    }
    // dfeat[n][s][c] = dropout1( sum_j dpre[n][j] * W1[j][c*S+s] )
    if (a.dfeat != nullptr)
        for (int i = tid; i < a.B * F; i += nt) {
            const int n = i / F, r = i % F;
            const int s = r / a.C, c = r % a.C;
            const int lf = c * a.S + s;
            float g = 0.f;
            for (int j = 0; j < a.Hd; ++j) g = fmaf(dh[n * a.Hd + j], sW1[j * FP + lf], g);
            if (a.p_drop > 0.f) g = (uniform01(a.seed, 1u, (uint32_t)(n * F + lf)) >= a.p_drop) ? g * keep_scale : 0.f;
            a.dfeat[i] = __float2bfloat16(g);
        }
}

}  // namespace nidt
using namespace nidt;

static int head_launch(const void* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                       float* gW1, float* gb1, float* gW2, float* gb2, void* dfeat, float* loss_acc, double* eval_counters,
                       float* logits_out, int B, int S, int C, int Hd, float p_drop, uint64_t seed, int train,
                       int double_sigmoid, const nidt_dyn_t* dyn, void* stream);
NIDT_API int nidt_head_fwd_bwd(const void* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                               float* gW1, float* gb1, float* gW2, float* gb2, void* dfeat, float* loss_acc, double* eval_counters,
                               float* logits_out, int B, int S, int C, int Hd, float p_drop, uint64_t seed, int train,
                               int double_sigmoid, void* stream) {
    return head_launch(feat, y, W1, b1, W2, b2, gW1, gb1, gW2, gb2, dfeat, loss_acc, eval_counters, logits_out, B, S, C, Hd, p_drop, seed, train,
                       double_sigmoid, nullptr, stream);
}
// seed and the loss accumulator (train) / eval counters (eval) are read from *dyn (DEVICE memory) by the kernel
NIDT_API int nidt_head_fwd_bwd_dyn(const void* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                                   float* gW1, float* gb1, float* gW2, float* gb2, void* dfeat, float* logits_out, int B, int S, int C, int Hd,
                                   float p_drop, int train, int double_sigmoid, const nidt_dyn_t* dyn, void* stream) {
    NIDT_REQUIRE(dyn != nullptr, "nidt_head_fwd_bwd_dyn: dyn is null");
    return head_launch(feat, y, W1, b1, W2, b2, gW1, gb1, gW2, gb2, dfeat, nullptr, nullptr, logits_out, B, S, C, Hd, p_drop, 0, train,
                       double_sigmoid, dyn, stream);
}
static int head_launch(const void* feat, const float* y, const float* W1, const float* b1, const float* W2, const float* b2,
                       float* gW1, float* gb1, float* gW2, float* gb2, void* dfeat, float* loss_acc, double* eval_counters,
                       float* logits_out, int B, int S, int C, int Hd, float p_drop, uint64_t seed, int train,
                       int double_sigmoid, const nidt_dyn_t* dyn, void* stream) {
    HeadArgs a{(const __nv_bfloat16*)feat, y, W1, b1, W2, b2, gW1, gb1, gW2, gb2, (__nv_bfloat16*)dfeat, loss_acc, eval_counters,
               logits_out, B, S, C, Hd, p_drop, seed, train, double_sigmoid, dyn};
    const size_t smem = ((size_t)B * S * C + 2 * (size_t)B * Hd + 2 * (size_t)B + (size_t)Hd * (S * C + 1)) * sizeof(float);
    NIDT_REQUIRE(smem <= 220 * 1024, "nidt_head_fwd_bwd: B*F=%d with hidden %d does not fit one CTA", B * S * C, Hd);
    NIDT_REQUIRE(B <= 256, "nidt_head_fwd_bwd: batch %d > 256", B);
    NIDT_CUDA_CHECK(cudaFuncSetAttribute(head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    head_kernel<<<1, 1024, smem, (cudaStream_t)stream>>>(a);
    NIDT_LAUNCH_CHECK();
    return 0;
}
