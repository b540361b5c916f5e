// Flat-slab kernels: everything that the reference does as per-tensor Python loops over a
// state_dict is one launch over the flat parameter slab here.
//
//   nidt_masked_sgd_step     clip_grad_norm_(10) + SGD(momentum, wd) + `w *= mask`
//                            (ref: sailentgrads/my_model_trainer.py:209,224-231)
//   nidt_weighted_aggregate  w_g = sum_i (n_i/sum n) w_i [* mask]   (ref: sailentgrads_api.py:212-227)
//   nidt_saliency_accumulate scores += |g*w|                          (ref: snip.py:60-71)
//   nidt_global_topk_mask    sum -> /sum -> k-th largest -> mask      (ref: snip.py:80-116)
//   nidt_count_nonzero / nidt_hamming / nidt_sq_dist                  (ref: model_trainer.py:49-53, slim_util.py:10-19)
#include "common.cuh"
#include "../../include/nidt_b200.h"
#include "topk.cuh"
#include <cooperative_groups.h>
#include <stdarg.h>
#include <string.h>

static thread_local char g_err[1024] = "";
void nidt_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
NIDT_API const char* nidt_last_error() { return g_err; }
NIDT_API int nidt_abi_version() { return 1; }

namespace nidt {

constexpr int kFlatThreads = 256;
constexpr int kNormBlocks = 2 * kNumSMs;  // fixed => the partial-sum tree is launch-invariant

// ---------------------------------------------------------------------------- grad norm
__global__ void __launch_bounds__(kFlatThreads) sq_norm_partials_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ partials) {
    __shared__ float red[kFlatThreads / 32];
    const int64_t n4 = n >> 2;
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kFlatThreads) {
        float4 v = ld4(g + 4 * i);
        acc0 = fmaf(v.x, v.x, acc0); acc1 = fmaf(v.y, v.y, acc1);
        acc2 = fmaf(v.z, v.z, acc2); acc3 = fmaf(v.w, v.w, acc3);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (int64_t i = n4 << 2; i < n; ++i) acc0 = fmaf(g[i], g[i], acc0);
    float t = block_sum<kFlatThreads>((acc0 + acc1) + (acc2 + acc3), red);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

// --------------------------------------------------------------------------- masked SGD
__global__ void __launch_bounds__(kFlatThreads) masked_sgd_kernel(
    float* __restrict__ w, const float* __restrict__ g, float* __restrict__ mom, const float* __restrict__ mask,
    int64_t n, float lr, float mu, float wd, float clip, int first_step,
    const float* __restrict__ partials, int n_partials, float* __restrict__ norm_out, const nidt_dyn_t* __restrict__ dyn) {
    __shared__ float red[kFlatThreads / 32];
    if (dyn != nullptr) { lr = dyn->lr; mu = dyn->momentum; wd = dyn->wd; clip = dyn->clip_norm; first_step = dyn->first_step; }
    float coef = 1.f;
    if (partials != nullptr) {
        float p = 0.f;
        for (int i = threadIdx.x; i < n_partials; i += kFlatThreads) p += partials[i];
        const float total = sqrtf(block_sum<kFlatThreads>(p, red));
        if (clip > 0.f) coef = fminf(clip / (total + 1e-6f), 1.f);
        if (blockIdx.x == 0 && threadIdx.x == 0 && norm_out) *norm_out = total;
    }
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kFlatThreads) {
        float4 wv = ld4(w + 4 * i);
        float4 gv = ld4(g + 4 * i);
        float4 d;
        d.x = fmaf(wd, wv.x, gv.x * coef); d.y = fmaf(wd, wv.y, gv.y * coef);
        d.z = fmaf(wd, wv.z, gv.z * coef); d.w = fmaf(wd, wv.w, gv.w * coef);
        if (mom != nullptr) {
            if (!first_step) {
                float4 mv = ld4(mom + 4 * i);
                d.x = fmaf(mu, mv.x, d.x); d.y = fmaf(mu, mv.y, d.y);
                d.z = fmaf(mu, mv.z, d.z); d.w = fmaf(mu, mv.w, d.w);
            }
            st4(mom + 4 * i, d);
        }
        wv.x = fmaf(-lr, d.x, wv.x); wv.y = fmaf(-lr, d.y, wv.y);
        wv.z = fmaf(-lr, d.z, wv.z); wv.w = fmaf(-lr, d.w, wv.w);
        if (mask != nullptr) {
            float4 mk = ld4(mask + 4 * i);
            wv.x *= mk.x; wv.y *= mk.y; wv.z *= mk.z; wv.w *= mk.w;
        }
        st4(w + 4 * i, wv);
    }
}

// -------------------------------------------------------------------------- aggregation
constexpr int kMaxAggSrc = 64;
struct AggArgs {
    const float* src[kMaxAggSrc];
    float wt[kMaxAggSrc];
    int k;
};

// out = sum_i wt[i]*src[i] in list order, multiply and add kept separate (no FMA contraction)
// so that the result is bit-identical to the reference's Python loop; optional `accumulate`
// continues a sum that was split over several launches (> kMaxAggSrc clients).
__global__ void __launch_bounds__(kFlatThreads) weighted_aggregate_kernel(
    float* __restrict__ out, const AggArgs a, const float* __restrict__ mask, int64_t n_mask, int64_t n, int accumulate, int last) {
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kFlatThreads) {
        float4 acc;
        int s = 0;
        if (accumulate) {
            acc = ld4(out + 4 * i);
        } else {
            float4 v = ld4(a.src[0] + 4 * i);
            const float wt = a.wt[0];
            acc.x = __fmul_rn(v.x, wt); acc.y = __fmul_rn(v.y, wt); acc.z = __fmul_rn(v.z, wt); acc.w = __fmul_rn(v.w, wt);
            s = 1;
        }
#pragma unroll 4
        for (; s < a.k; ++s) {
            float4 v = ld4(a.src[s] + 4 * i);
            const float wt = a.wt[s];
            acc.x = __fadd_rn(acc.x, __fmul_rn(v.x, wt)); acc.y = __fadd_rn(acc.y, __fmul_rn(v.y, wt));
            acc.z = __fadd_rn(acc.z, __fmul_rn(v.z, wt)); acc.w = __fadd_rn(acc.w, __fmul_rn(v.w, wt));
        }
        if (last && mask != nullptr && 4 * i < n_mask) {
            float4 mk = ld4(mask + 4 * i);
            acc.x *= mk.x; acc.y *= mk.y; acc.z *= mk.z; acc.w *= mk.w;
        }
        st4(out + 4 * i, acc);
    }
}

// ------------------------------------------------------------------------------ saliency
__global__ void __launch_bounds__(kFlatThreads) saliency_accumulate_kernel(
    float* __restrict__ scores, const float* __restrict__ g, const float* __restrict__ w, int64_t n, float scale) {
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n4; i += (int64_t)gridDim.x * kFlatThreads) {
        float4 s = ld4(scores + 4 * i), gv = ld4(g + 4 * i), wv = ld4(w + 4 * i);
        s.x = fmaf(scale, fabsf(gv.x * wv.x), s.x); s.y = fmaf(scale, fabsf(gv.y * wv.y), s.y);
        s.z = fmaf(scale, fabsf(gv.z * wv.z), s.z); s.w = fmaf(scale, fabsf(gv.w * wv.w), s.w);
        st4(scores + 4 * i, s);
    }
}

// ------------------------------------------------------------------ global top-k -> mask
// One cooperative launch: (1) sum of the prunable scores (fixed-order block partials), (2) four
// 8-bit radix-select passes over the fp32 bit pattern of q = s/sum (q >= 0, so the unsigned bit
// pattern is order preserving) to find the k-th largest q exactly, (3) mask = (q >= thr) on
// prunable entries, 1 on other real parameters, 0 on slab padding.
// `flags[i]`: 0 = padding, 1 = real but never pruned (bias / BN affine), 2 = prunable weight.
__global__ void __launch_bounds__(kTopkThreads) global_topk_mask_kernel(
    const float* __restrict__ scores, const unsigned char* __restrict__ flags, float* __restrict__ mask,
    int64_t n, unsigned long long k, TopkScratch* __restrict__ sc) {
    cg::grid_group grid = cg::this_grid();
    topk_mask_body(grid, scores, flags, mask, n, k, sc);
}

// ----------------------------------------------------------------------------- statistics
// mode 0: count(a != 0); mode 1: count((a != 0) xor (b != 0)); result added to *out (u64)
__global__ void __launch_bounds__(kFlatThreads) count_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n, int mode,
                                                             unsigned long long* __restrict__ out) {
    unsigned long long c = 0ull;
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kFlatThreads) {
        const bool x = a[i] != 0.f;
        c += mode == 0 ? (unsigned long long)x : (unsigned long long)(x != (b[i] != 0.f));
    }
    c = warp_sum_u64(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

__global__ void __launch_bounds__(kFlatThreads) sq_dist_partials_kernel(const float* __restrict__ a, const float* __restrict__ b, int64_t n,
                                                                        float* __restrict__ partials) {
    __shared__ float red[kFlatThreads / 32];
    float acc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kFlatThreads) {
        const float d = a[i] - b[i];
        acc = fmaf(d, d, acc);
    }
    float t = block_sum<kFlatThreads>(acc, red);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

__global__ void __launch_bounds__(kFlatThreads) sum_partials_kernel(const float* __restrict__ partials, int n, float* __restrict__ out, int take_sqrt) {
    __shared__ float red[kFlatThreads / 32];
    float p = 0.f;
    for (int i = threadIdx.x; i < n; i += kFlatThreads) p += partials[i];
    float t = block_sum<kFlatThreads>(p, red);
    if (threadIdx.x == 0) *out = take_sqrt ? sqrtf(t) : t;
}

// Ditto proximal pull:  w -= lr*lamda*(w - w_ref)
__global__ void __launch_bounds__(kFlatThreads) prox_pull_kernel(float* __restrict__ w, const float* __restrict__ ref, int64_t n, float coef) {
    for (int64_t i = (int64_t)blockIdx.x * kFlatThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kFlatThreads)
        w[i] = fmaf(-coef, w[i] - ref[i], w[i]);
}

static inline int flat_grid(int64_t n_items) {
    int64_t b = ceil_div64(n_items, kFlatThreads);
    const int64_t cap = (int64_t)kNumSMs * 8;
    return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace nidt

using namespace nidt;

NIDT_API int nidt_norm_scratch_floats() { return kNormBlocks + 4; }
NIDT_API int nidt_topk_scratch_bytes() { return (int)sizeof(TopkScratch); }
NIDT_API int nidt_max_agg_sources() { return kMaxAggSrc; }

// scratch: >= nidt_norm_scratch_floats() floats.  norm_out may be null.
static int masked_sgd_step(float* w, const float* g, float* mom, const float* mask, int64_t n, float lr, float momentum, float wd,
                           float clip_norm, int first_step, float* scratch, float* norm_out, const nidt_dyn_t* dyn, void* stream);
NIDT_API int nidt_masked_sgd_step(float* w, const float* g, float* mom, const float* mask, int64_t n,
                                    float lr, float momentum, float wd, float clip_norm, int first_step,
                                    float* scratch, float* norm_out, void* stream) {
    return masked_sgd_step(w, g, mom, mask, n, lr, momentum, wd, clip_norm, first_step, scratch, norm_out, nullptr, stream);
}
// dyn (DEVICE pointer): lr / momentum / wd / clip / first_step are read from it by the kernel; `mom` must already be null when no
// momentum buffer is to be used (the host cannot look at dyn->momentum).
NIDT_API int nidt_masked_sgd_step_dyn(float* w, const float* g, float* mom, const float* mask, int64_t n, float* scratch, const nidt_dyn_t* dyn,
                                        void* stream) {
    NIDT_REQUIRE(dyn != nullptr, "nidt_masked_sgd_step_dyn: dyn is null");
    return masked_sgd_step(w, g, mom, mask, n, 0.f, mom != nullptr ? 1.f : 0.f, 0.f, 0.f, 0, scratch, nullptr, dyn, stream);
}
static int masked_sgd_step(float* w, const float* g, float* mom, const float* mask, int64_t n, float lr, float momentum, float wd,
                           float clip_norm, int first_step, float* scratch, float* norm_out, const nidt_dyn_t* dyn, void* stream) {
    NIDT_REQUIRE(n % 4 == 0, "nidt_masked_sgd_step: n=%lld must be a multiple of 4 (slab alignment)", (long long)n);
    NIDT_REQUIRE(scratch != nullptr, "nidt_masked_sgd_step: scratch is null");
    cudaStream_t s = (cudaStream_t)stream;
    sq_norm_partials_kernel<<<kNormBlocks, kFlatThreads, 0, s>>>(g, n, scratch);
    NIDT_LAUNCH_CHECK();
    masked_sgd_kernel<<<flat_grid(n / 4), kFlatThreads, 0, s>>>(w, g, momentum != 0.f ? mom : nullptr, mask, n, lr, momentum, wd,
                                                                  clip_norm, first_step, scratch, kNormBlocks, norm_out, dyn);
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_grad_norm(const float* g, int64_t n, float* scratch, float* norm_out, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    sq_norm_partials_kernel<<<kNormBlocks, kFlatThreads, 0, s>>>(g, n, scratch);
    NIDT_LAUNCH_CHECK();
    sum_partials_kernel<<<1, kFlatThreads, 0, s>>>(scratch, kNormBlocks, norm_out, 1);
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_weighted_aggregate(float* out, const float* const* srcs, const float* weights, int k,
                                       const float* mask, int64_t n_mask, int64_t n, void* stream) {
    NIDT_REQUIRE(k >= 1, "nidt_weighted_aggregate: need at least one source");
    NIDT_REQUIRE(n % 4 == 0 && n_mask % 4 == 0, "nidt_weighted_aggregate: n and n_mask must be multiples of 4");
    cudaStream_t s = (cudaStream_t)stream;
    for (int base = 0; base < k; base += kMaxAggSrc) {
        AggArgs a;
        a.k = k - base < kMaxAggSrc ? k - base : kMaxAggSrc;
        for (int i = 0; i < a.k; ++i) { a.src[i] = srcs[base + i]; a.wt[i] = weights[base + i]; }
        const int last = base + a.k >= k;
        weighted_aggregate_kernel<<<flat_grid(n / 4), kFlatThreads, 0, s>>>(out, a, mask, n_mask, n, base > 0, last);
        NIDT_LAUNCH_CHECK();
    }
    return 0;
}

NIDT_API int nidt_saliency_accumulate(float* scores, const float* g, const float* w, int64_t n, float scale, void* stream) {
    NIDT_REQUIRE(n % 4 == 0, "nidt_saliency_accumulate: n must be a multiple of 4");
    saliency_accumulate_kernel<<<flat_grid(n / 4), kFlatThreads, 0, (cudaStream_t)stream>>>(scores, g, w, n, scale);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// scratch: >= nidt_topk_scratch_bytes() bytes of device memory (contents ignored on entry).
// On return (stream-ordered) scratch holds {.., threshold, norm, kept} -- see TopkScratch.
NIDT_API int nidt_global_topk_mask(const float* scores, const unsigned char* flags, float* mask, int64_t n,
                                     unsigned long long k, void* scratch, void* stream) {
    NIDT_REQUIRE(scratch != nullptr, "nidt_global_topk_mask: scratch is null");
    int dev = 0, sms = 0, per_sm = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    NIDT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, global_topk_mask_kernel, kTopkThreads, 0));
    NIDT_REQUIRE(per_sm >= 1, "nidt_global_topk_mask: kernel does not fit on an SM");
    int grid = sms * (per_sm > 2 ? 2 : per_sm);
    if (grid > 2 * kNumSMs) grid = 2 * kNumSMs;
    TopkScratch* sc = (TopkScratch*)scratch;
    void* args[] = {(void*)&scores, (void*)&flags, (void*)&mask, (void*)&n, (void*)&k, (void*)&sc};
    NIDT_CUDA_CHECK(cudaLaunchCooperativeKernel((void*)global_topk_mask_kernel, dim3(grid), dim3(kTopkThreads), args, 0, (cudaStream_t)stream));
    return 0;
}

NIDT_API int nidt_topk_read_result(const void* scratch, float* threshold, float* norm, unsigned int* kept, void* stream) {
    TopkScratch h;
    NIDT_CUDA_CHECK(cudaMemcpyAsync(&h, scratch, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    NIDT_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    if (threshold) *threshold = h.threshold;
    if (norm) *norm = h.norm;
    if (kept) *kept = h.kept;
    return 0;
}

// *out (device u64) must be zeroed by the caller (or hold a running total to add to).
NIDT_API int nidt_count_nonzero(const float* a, int64_t n, unsigned long long* out, void* stream) {
    count_kernel<<<flat_grid(n), kFlatThreads, 0, (cudaStream_t)stream>>>(a, nullptr, n, 0, out);
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_hamming(const float* a, const float* b, int64_t n, unsigned long long* out, void* stream) {
    count_kernel<<<flat_grid(n), kFlatThreads, 0, (cudaStream_t)stream>>>(a, b, n, 1, out);
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_sq_dist(const float* a, const float* b, int64_t n, float* scratch, float* out, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    sq_dist_partials_kernel<<<kNormBlocks, kFlatThreads, 0, s>>>(a, b, n, scratch);
    NIDT_LAUNCH_CHECK();
    sum_partials_kernel<<<1, kFlatThreads, 0, s>>>(scratch, kNormBlocks, out, 0);
    NIDT_LAUNCH_CHECK();
    return 0;
}

NIDT_API int nidt_prox_pull(float* w, const float* ref, int64_t n, float coef, void* stream) {
    prox_pull_kernel<<<flat_grid(n), kFlatThreads, 0, (cudaStream_t)stream>>>(w, ref, n, coef);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------- batch gather of 8-bit volumes
// dst[j][0..row_bytes) = src[idx[j]][0..row_bytes) for rows of ANY byte length (an ABCD volume is 2 122 945 bytes, so neither
// side is 16-byte aligned in general): every thread writes one aligned 16-byte vector of dst assembled from the two aligned
// source vectors that cover it (byte funnel shift); the unaligned head and tail of the whole destination are written bytewise.
// Replaces torch.index_select on uint8 (~0.7 TB/s) when batches are assembled from a device-resident dataset.
namespace nidt {
__device__ __forceinline__ uint32_t funnel_bytes(uint32_t lo, uint32_t hi, uint32_t byte_shift) {
    return __funnelshift_r(lo, hi, byte_shift * 8);
}
constexpr int kGatherArgRows = 64;
struct RowArgs { int64_t v[kGatherArgRows]; __device__ __forceinline__ int64_t operator[](int j) const { return v[j]; } };
template <class Idx>
__device__ __forceinline__ void gather_rows_body(const uint8_t* __restrict__ src, const Idx& idx, uint8_t* __restrict__ dst, int rows, int64_t row_bytes) {
    const int64_t total = (int64_t)rows * row_bytes;
    const uintptr_t d0 = reinterpret_cast<uintptr_t>(dst);
    const int64_t head = (int64_t)((16 - (d0 & 15)) & 15) < total ? (int64_t)((16 - (d0 & 15)) & 15) : total;
    const int64_t nvec = (total - head) >> 4;
    const int64_t tail0 = head + (nvec << 4);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    for (int64_t v = tid; v < nvec; v += nth) {
        const int64_t o = head + (v << 4);              // dst byte offset of this vector
        const int j = (int)(o / row_bytes);
        const int64_t within = o - (int64_t)j * row_bytes;
        uint32_t out[4];
        if (within + 16 <= row_bytes) {                 // the vector lies inside one row: two aligned source vectors cover it
            const uint8_t* s = src + idx[j] * row_bytes + within;
            const uintptr_t sa = reinterpret_cast<uintptr_t>(s);
            const uint4 a = *reinterpret_cast<const uint4*>(sa & ~(uintptr_t)15);
            const uint32_t sh = (uint32_t)(sa & 15);
            uint32_t w[8] = {a.x, a.y, a.z, a.w, 0u, 0u, 0u, 0u};
            if (sh) {
                const uint4 b = *reinterpret_cast<const uint4*>((sa & ~(uintptr_t)15) + 16);
                w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
            }
            const uint32_t ws = sh >> 2, bs = sh & 3;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint32_t lo = 0, hi = 0;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (ws == (uint32_t)q) { lo = w[k + q]; hi = w[k + q + 1]; }
                out[k] = funnel_bytes(lo, hi, bs);
            }
        } else {                                        // straddles two rows: bytewise
            uint8_t* ob = reinterpret_cast<uint8_t*>(out);
            for (int b = 0; b < 16; ++b) {
                const int64_t ob_off = o + b;
                const int jj = (int)(ob_off / row_bytes);
                ob[b] = src[idx[jj] * row_bytes + (ob_off - (int64_t)jj * row_bytes)];
            }
        }
        *reinterpret_cast<uint4*>(dst + o) = make_uint4(out[0], out[1], out[2], out[3]);
    }
    for (int64_t o = tid; o < head + (total - tail0); o += nth) {      // unaligned head and tail bytes of the destination
        const int64_t off = o < head ? o : tail0 + (o - head);
        const int j = (int)(off / row_bytes);
        dst[off] = src[idx[j] * row_bytes + (off - (int64_t)j * row_bytes)];
    }
}
__global__ void __launch_bounds__(256) gather_rows_u8_kernel(const uint8_t* __restrict__ src, const int64_t* __restrict__ idx, uint8_t* __restrict__ dst,
                                                             int rows, int64_t row_bytes) {
    gather_rows_body(src, idx, dst, rows, row_bytes);
}
// Row numbers travel in the kernel parameters (no index buffer on the device, no host->device copy of it); the labels of the
// batch are gathered by the same launch.
__global__ void __launch_bounds__(256) gather_rows_u8_args_kernel(const uint8_t* __restrict__ src, const __grid_constant__ RowArgs idx, uint8_t* __restrict__ dst,
                                                                  int rows, int64_t row_bytes, const float* __restrict__ ysrc, float* __restrict__ ydst) {
    gather_rows_body(src, idx, dst, rows, row_bytes);
    if (ysrc != nullptr && blockIdx.x == 0 && threadIdx.x < rows) ydst[threadIdx.x] = ysrc[idx[threadIdx.x]];
}
}  // namespace nidt

// idx: int64 device array of `rows` source row numbers.  The source allocation must be readable up to the next 16-byte boundary
// past its last row (true for any cudaMalloc'ed / PyTorch buffer: allocations are padded to at least 256 bytes).
static int gather_rows_launch(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes, int max_blocks, void* stream);
NIDT_API int nidt_gather_rows_u8(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes, void* stream) {
    return gather_rows_launch(src, idx, dst, rows, row_bytes, kNumSMs * 16, stream);
}
// Same gather with a capped grid: src / idx may be PINNED HOST memory (read over PCIe by the kernel itself), so a batch is
// staged by a few CTAs next to the compute kernels instead of by the copy engine.
NIDT_API int nidt_gather_rows_u8_capped(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes, int max_blocks, void* stream) {
    return gather_rows_launch(src, idx, dst, rows, row_bytes, max_blocks, stream);
}
static int gather_rows_launch(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes, int max_blocks, void* stream) {
    NIDT_REQUIRE(rows >= 0 && row_bytes >= 16 && max_blocks >= 1, "nidt_gather_rows_u8: rows=%d row_bytes=%lld", rows, (long long)row_bytes);
    if (rows == 0) return 0;
    const int64_t nvec = ((int64_t)rows * row_bytes) >> 4;
    int64_t grid = (nvec + 255) / 256;
    if (grid > (int64_t)max_blocks) grid = (int64_t)max_blocks;
    if (grid < 1) grid = 1;
    gather_rows_u8_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)src, idx, (uint8_t*)dst, rows, row_bytes);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// Gather with the row numbers passed by value (rows <= 64; `idx` is a HOST array) + optional fp32 label gather ydst[j] = ysrc[idx[j]].
NIDT_API int nidt_gather_rows_u8_args(const void* src, const int64_t* idx, void* dst, int rows, int64_t row_bytes,
                                      const float* ysrc, float* ydst, void* stream) {
    NIDT_REQUIRE(rows >= 0 && rows <= kGatherArgRows && row_bytes >= 16, "nidt_gather_rows_u8_args: rows=%d (max %d) row_bytes=%lld", rows,
                 kGatherArgRows, (long long)row_bytes);
    if (rows == 0) return 0;
    RowArgs a;
    for (int j = 0; j < kGatherArgRows; ++j) a.v[j] = j < rows ? idx[j] : 0;
    const int64_t nvec = ((int64_t)rows * row_bytes) >> 4;
    int64_t grid = (nvec + 255) / 256;
    if (grid > (int64_t)kNumSMs * 16) grid = (int64_t)kNumSMs * 16;
    if (grid < 1) grid = 1;
    gather_rows_u8_args_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)src, a, (uint8_t*)dst, rows, row_bytes, ysrc, ydst);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------- zero-run-skipping volume transfer
// Masked grey-matter maps are mostly background: outside the brain every voxel is 0 (Preprocess_ABCD.ipynb applies a group-mean
// mask), so each image row is [zeros | span | zeros].  The host keeps, per volume, the spans packed back to back plus a 6-byte
// table entry per row (start, length, offset of the span in the packed volume); only those cross PCIe, and this kernel rebuilds the
// dense 8-bit volume on the device: one warp per image row, every byte of the row written (span bytes or 0).
namespace nidt {
constexpr int kExpandVolumes = 32;
struct ExpandArgs {
    long long packed_off[kExpandVolumes];      // byte offset of every volume's packed spans in `packed`
};
__global__ void __launch_bounds__(256) expand_sparse_rows_kernel(const uint8_t* __restrict__ packed, const __grid_constant__ ExpandArgs a,
                                                                 const uint8_t* __restrict__ tab /* [vol][row][2] */, const uint32_t* __restrict__ off /* [vol][row] */,
                                                                 uint8_t* __restrict__ dst, int n_vol, int rows, int W) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t total = (int64_t)n_vol * rows;
    for (int64_t r = warp; r < total; r += n_warps) {
        const int v = (int)(r / rows);
        const unsigned start = tab[2 * r], len = tab[2 * r + 1];
        const uint8_t* src = packed + a.packed_off[v] + off[r];
        uint8_t* out = dst + r * W;
        for (int c = lane; c < W; c += 32) {
            const unsigned k = (unsigned)c - start;
            out[c] = k < len ? src[k] : (uint8_t)0;
        }
    }
}
}  // namespace nidt
// packed / tab / off / dst: DEVICE buffers (the packed spans and tables of `n_vol` <= 32 volumes as uploaded); packed_off: HOST array
NIDT_API int nidt_expand_sparse_rows_u8(const void* packed, const long long* packed_off, const void* tab, const void* off, void* dst, int n_vol,
                                        int rows, int W, void* stream) {
    NIDT_REQUIRE(n_vol >= 0 && n_vol <= nidt::kExpandVolumes && rows >= 1 && W >= 1 && W <= 255, "nidt_expand_sparse_rows_u8: n_vol=%d rows=%d W=%d", n_vol, rows, W);
    if (n_vol == 0) return 0;
    nidt::ExpandArgs a;
    for (int i = 0; i < nidt::kExpandVolumes; ++i) a.packed_off[i] = i < n_vol ? packed_off[i] : 0;
    const int64_t warps = (int64_t)n_vol * rows;
    int64_t grid = (warps + 7) / 8;
    if (grid > nidt::kNumSMs * 4) grid = nidt::kNumSMs * 4;       // a few CTAs per SM: it runs next to the training kernels
    nidt::expand_sparse_rows_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>((const uint8_t*)packed, a, (const uint8_t*)tab, (const uint32_t*)off,
                                                                                (uint8_t*)dst, n_vol, rows, W);
    NIDT_LAUNCH_CHECK();
    return 0;
}

// dst[j] = host_base[rows[j]] (row_bytes each) as asynchronous copies on `stream`, issued from native code.  Runs of consecutive
// row numbers (a site's subjects are adjacent in the store) become ONE copy each, so a sorted batch costs a handful of large
// transfers instead of one per sample.  host_base should be pinned; rows is a HOST array.  *n_copies (optional) = copies issued.
static int copy_rows_h2d(const void* host_base, const int64_t* rows, int n, int64_t row_bytes, void* dst, void* stream, int* n_copies) {
    NIDT_REQUIRE(n >= 0 && row_bytes > 0, "nidt_copy_rows_h2d: n=%d row_bytes=%lld", n, (long long)row_bytes);
    int copies = 0;
    for (int j = 0; j < n;) {
        int e = j + 1;
        while (e < n && rows[e] == rows[e - 1] + 1) ++e;
        NIDT_CUDA_CHECK(cudaMemcpyAsync((uint8_t*)dst + (size_t)j * row_bytes, (const uint8_t*)host_base + (size_t)rows[j] * row_bytes,
                                        (size_t)(e - j) * row_bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
        ++copies;
        j = e;
    }
    if (n_copies) *n_copies = copies;
    return 0;
}
NIDT_API int nidt_copy_rows_h2d(const void* host_base, const int64_t* rows, int n, int64_t row_bytes, void* dst, void* stream) {
    return copy_rows_h2d(host_base, rows, n, row_bytes, dst, stream, nullptr);
}
NIDT_API int nidt_copy_rows_h2d_counted(const void* host_base, const int64_t* rows, int n, int64_t row_bytes, void* dst, void* stream, int* n_copies) {
    return copy_rows_h2d(host_base, rows, n, row_bytes, dst, stream, n_copies);
}

// ------------------------------------------------------------------------------------------------ host side of the zero-run-skipping uploads
// The span table of a volume cache (runtime/staging.py::SparsePack): for every image row of W bytes the first non-zero byte and the
// length up to the last one, and the row's offset among its volume's packed spans.  Plain host code on a few threads: the numpy
// statement of the same scan takes seconds per GB, which every rank pays at start-up.
#include <thread>
#include <vector>
namespace {
template <class F>
void parallel_volumes(int64_t n_vol, int threads, F f) {
    if (threads < 1) threads = 1;
    if (threads > 64) threads = 64;
    if ((int64_t)threads > n_vol) threads = (int)(n_vol > 0 ? n_vol : 1);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([=] { for (int64_t v = t; v < n_vol; v += threads) f(v); });
    for (auto& th : pool) th.join();
}
}  // namespace

// tab: [n_vol][rows][2] bytes (start, length); off: [n_vol][rows] uint32; vol_bytes: [n_vol] int64 (packed bytes of each volume)
NIDT_API int nidt_sparse_scan_rows(const unsigned char* cache, int64_t n_vol, int rows, int W, unsigned char* tab, unsigned int* off,
                                   int64_t* vol_bytes, int threads) {
    NIDT_REQUIRE(cache != nullptr && tab != nullptr && off != nullptr && vol_bytes != nullptr, "nidt_sparse_scan_rows: null argument");
    NIDT_REQUIRE(n_vol >= 0 && rows >= 1 && W >= 1 && W <= 255, "nidt_sparse_scan_rows: n_vol=%lld rows=%d W=%d (W <= 255)", (long long)n_vol, rows, W);
    parallel_volumes(n_vol, threads, [=](int64_t v) {
        const unsigned char* p = cache + (size_t)v * rows * W;
        unsigned char* tb = tab + (size_t)v * rows * 2;
        unsigned int* of = off + (size_t)v * rows;
        unsigned int acc = 0;
        for (int r = 0; r < rows; ++r, p += W) {
            int a = 0, b = W - 1;
            while (a < W && p[a] == 0) ++a;
            int len = 0;
            if (a < W) {
                while (p[b] == 0) --b;
                len = b - a + 1;
            } else {
                a = 0;
            }
            tb[2 * r] = (unsigned char)a;
            tb[2 * r + 1] = (unsigned char)len;
            of[r] = acc;
            acc += (unsigned int)len;
        }
        vol_bytes[v] = (int64_t)acc;
    });
    return 0;
}

// packed[vol_off[v] + off[v][r] ...] = the span of row r of volume v
NIDT_API int nidt_sparse_copy_rows(const unsigned char* cache, int64_t n_vol, int rows, int W, const unsigned char* tab, const unsigned int* off,
                                   const int64_t* vol_off, unsigned char* packed, int threads) {
    NIDT_REQUIRE(cache != nullptr && tab != nullptr && off != nullptr && vol_off != nullptr && (packed != nullptr || n_vol == 0),
                 "nidt_sparse_copy_rows: null argument");
    parallel_volumes(n_vol, threads, [=](int64_t v) {
        const unsigned char* p = cache + (size_t)v * rows * W;
        const unsigned char* tb = tab + (size_t)v * rows * 2;
        const unsigned int* of = off + (size_t)v * rows;
        unsigned char* dst = packed + vol_off[v];
        for (int r = 0; r < rows; ++r, p += W)
            if (tb[2 * r + 1]) memcpy(dst + of[r], p + tb[2 * r], tb[2 * r + 1]);
    });
    return 0;
}
