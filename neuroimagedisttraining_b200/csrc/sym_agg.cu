// Cross-GPU aggregation fused with its collective (north-star kernels W1 / W2 of SURVEY §8).
//
// Every rank holds, in NVLink-mapped symmetric memory, `partial` = the sample-weighted sum of the
// models of its own sampled clients (one flat launch, flat_ops.cu).  ONE kernel per rank then does
//     reduce-scatter : rank r sums chunk r of all ranks' partials with P2P loads (fixed rank order ->
//                      every rank computes bit-identical values), or with multimem.ld_reduce (NVLS,
//                      in-switch reduction) when a multicast address is available and requested;
//     mask           : multiplies the SalientGrads mask into the reduced chunk;
//     all-gather     : writes the chunk into every rank's `result` buffer with P2P stores (or one
//                      multimem.st), i.e. the new global model lands in all slabs at once;
//     write-back     : copies `result` into the caller's (non-symmetric) global slab.
// Cross-rank ordering uses release/acquire flags in symmetric memory with a monotonically increasing
// epoch (no resets), grid-wide ordering uses a cooperative launch.  No NCCL call, no extra passes.
// Reference: the CPU dict loop `_aggregate` (fedml_api/standalone/sailentgrads/sailentgrads_api.py:212-227)
// and, for the score reduction, snip.py:120-140.
//
// The same reduce+broadcast phase, followed by the radix-select of flat_ops.cu, is the fused
// "saliency all-reduce + global top-k + mask emission" (nidt_sym_saliency_topk_mask).
#include "common.cuh"
#include "topk.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>

namespace nidt {

constexpr int kMaxRanks = 16;
constexpr int kSymThreads = 512;

struct SymPeers {
    float* partial[kMaxRanks];      // peer-mapped base of every rank's partial buffer
    float* result[kMaxRanks];       // peer-mapped base of every rank's result buffer
    uint32_t* flags[kMaxRanks];     // peer-mapped flag pads: flags[dst][slot*kMaxRanks + src]
    float* mc_partial;              // multicast address of `partial` (nullptr = unavailable)
    float* mc_result;               // multicast address of `result`
    int rank, world;
    long long timeout_cycles;       // a peer that has not arrived after this many SM clocks ends the kernel in a trap (hang detector only:
                                    // the host aligns the ranks before the launch, see parallel/symmetric.py)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// All ranks arrive, then all ranks leave.  Block 0 does the signalling; `value` must grow by one per
// barrier over the lifetime of the flag pad.  A peer that never arrives ends in a trap, not a hang.
__device__ __forceinline__ void cross_rank_barrier(cg::grid_group& grid, const SymPeers& sp, int slot, uint32_t value, bool order_prior_writes = true) {
    if (order_prior_writes) {      // entry barrier of a kernel: the data it publishes was written by earlier launches on this stream
        __threadfence_system();
        grid.sync();
    }
    if (blockIdx.x == 0 && threadIdx.x < sp.world) {
        const int peer = threadIdx.x;
        st_release_sys(sp.flags[peer] + slot * kMaxRanks + sp.rank, value);
        const uint32_t* mine = sp.flags[sp.rank] + slot * kMaxRanks + peer;
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(mine) - value) < 0) {
            if (clock64() - t0 > sp.timeout_cycles) {
                printf("nidt: cross-rank barrier timed out (rank %d waiting for %d, value %u)\n", sp.rank, peer, value);
                __trap();
            }
        }
    }
    grid.sync();
}

__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st_f32x4(float* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// reduce chunk `rank` over all partials (+mask) and broadcast it into every result buffer.
// kWorld (2/4/8/16 >= world) bounds the registers; two 16-byte vectors per thread and iteration are in flight per peer.
template <int kWorld>
__device__ __forceinline__ void reduce_mask_broadcast_p2p(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t n) {
    const int64_t n4 = n >> 2;
    const int64_t per = (n4 + sp.world - 1) / sp.world;
    const int64_t lo = per * sp.rank, hi = min(n4, lo + per);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    constexpr int kVec = kWorld <= 4 ? 2 : 1;       // keep <= 64 registers so two CTAs fit per SM
    for (int64_t i0 = lo + tid; i0 < hi; i0 += kVec * nth) {
        const int64_t i1 = i0 + nth;
        const bool two = kVec == 2 && i1 < hi;
        float4 a[kWorld], b[kWorld];
#pragma unroll
        for (int p = 0; p < kWorld; ++p)
            if (p < sp.world) {
                a[p] = ld4(sp.partial[p] + 4 * i0);
                if (two) b[p] = ld4(sp.partial[p] + 4 * i1);
            }
        float4 s0 = a[0], s1 = two ? b[0] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 1; p < kWorld; ++p)
            if (p < sp.world) {
                s0.x += a[p].x; s0.y += a[p].y; s0.z += a[p].z; s0.w += a[p].w;
                if (two) { s1.x += b[p].x; s1.y += b[p].y; s1.z += b[p].z; s1.w += b[p].w; }
            }
        if (mask != nullptr) {
            if (4 * i0 < n_mask) { const float4 mk = ld4(mask + 4 * i0); s0.x *= mk.x; s0.y *= mk.y; s0.z *= mk.z; s0.w *= mk.w; }
            if (two && 4 * i1 < n_mask) { const float4 mk = ld4(mask + 4 * i1); s1.x *= mk.x; s1.y *= mk.y; s1.z *= mk.z; s1.w *= mk.w; }
        }
#pragma unroll
        for (int p = 0; p < kWorld; ++p)
            if (p < sp.world) {
                st4(sp.result[p] + 4 * i0, s0);
                if (two) st4(sp.result[p] + 4 * i1, s1);
            }
    }
}

__device__ __forceinline__ void reduce_mask_broadcast_multimem(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t n) {
    const int64_t n4 = n >> 2;
    const int64_t per = (n4 + sp.world - 1) / sp.world;
    const int64_t lo = per * sp.rank, hi = min(n4, lo + per);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    constexpr int U = 4;
    for (int64_t i = lo + tid; i < hi; i += U * nth) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (i + u * nth < hi) v[u] = multimem_ld_reduce_f32x4(sp.mc_partial + 4 * (i + u * nth));
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = i + u * nth;
            if (j < hi) {
                if (mask != nullptr && 4 * j < n_mask) { const float4 mk = ld4(mask + 4 * j); v[u].x *= mk.x; v[u].y *= mk.y; v[u].z *= mk.z; v[u].w *= mk.w; }
                multimem_st_f32x4(sp.mc_result + 4 * j, v[u]);
            }
        }
    }
}

// kWorld == 0 selects the NVLS multimem path
template <int kWorld>
__device__ __forceinline__ void reduce_mask_broadcast(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t n) {
    if constexpr (kWorld == 0) reduce_mask_broadcast_multimem(sp, mask, n_mask, n);
    else reduce_mask_broadcast_p2p<kWorld>(sp, mask, n_mask, n);
}

template <int kWorld>
__global__ void __launch_bounds__(kSymThreads, kWorld <= 8 ? 2 : 1) sym_agg_kernel(const SymPeers sp, const float* __restrict__ mask, int64_t n_mask, int64_t n,
                                                                                   float* __restrict__ out_local, uint32_t epoch) {
    cg::grid_group grid = cg::this_grid();
    cross_rank_barrier(grid, sp, 0, 2 * epoch + 1, false);   // every rank's partial is complete
    reduce_mask_broadcast<kWorld>(sp, mask, n_mask, n);
    cross_rank_barrier(grid, sp, 0, 2 * epoch + 2);          // every chunk has landed everywhere
    if (out_local != nullptr && out_local != sp.result[sp.rank]) {
        const float* src = sp.result[sp.rank];
        const int64_t n4 = n >> 2;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
            st4(out_local + 4 * i, ld4(src + 4 * i));
    }
}

// W2: cross-rank saliency reduction fused with the global top-k threshold and the mask emission.
// `partial` holds this rank's sum of its clients' (already 1/num_clients scaled) |g*w| scores.
template <int kWorld>
__global__ void __launch_bounds__(kSymThreads) sym_saliency_topk_kernel(const SymPeers sp, int64_t n, const unsigned char* __restrict__ flags,
                                                                        float* __restrict__ mask, unsigned long long k, TopkScratch* __restrict__ sc,
                                                                        float* __restrict__ scores_out, uint32_t epoch) {
    static_assert(kSymThreads == kTopkThreads, "block size is shared with the top-k body");
    cg::grid_group grid = cg::this_grid();
    cross_rank_barrier(grid, sp, 0, 2 * epoch + 1, false);
    reduce_mask_broadcast<kWorld>(sp, nullptr, 0, n);
    cross_rank_barrier(grid, sp, 0, 2 * epoch + 2);
    const float* scores = sp.result[sp.rank];
    topk_mask_body(grid, scores, flags, mask, n, k, sc);
    if (scores_out != nullptr && scores_out != scores)
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) scores_out[i] = scores[i];
}

// NIDT_SYM_TIMEOUT_S (default 300 s) at the nominal 1.965 GHz SM clock.
static long long sym_timeout_cycles() {
    static long long cached = 0;
    if (cached == 0) {
        double sec = 300.0;
        if (const char* e = getenv("NIDT_SYM_TIMEOUT_S")) { const double v = atof(e); if (v > 0.0) sec = v; }
        cached = (long long)(sec * 1.965e9);
    }
    return cached;
}

static int fill_peers(SymPeers& sp, void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial, void* mc_result,
                      int rank, int world) {
    NIDT_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "sym: bad rank %d / world %d", rank, world);
    for (int p = 0; p < kMaxRanks; ++p) {
        sp.partial[p] = p < world ? (float*)partial_ptrs[p] : nullptr;
        sp.result[p] = p < world ? (float*)result_ptrs[p] : nullptr;
        sp.flags[p] = p < world ? (uint32_t*)flag_ptrs[p] : nullptr;
    }
    sp.mc_partial = (float*)mc_partial;
    sp.mc_result = (float*)mc_result;
    sp.rank = rank;
    sp.world = world;
    sp.timeout_cycles = sym_timeout_cycles();
    return 0;
}

static const void* pick_agg_kernel(int world, int use_multimem) {
    if (use_multimem) return (const void*)sym_agg_kernel<0>;
    if (world <= 2) return (const void*)sym_agg_kernel<2>;
    if (world <= 4) return (const void*)sym_agg_kernel<4>;
    if (world <= 8) return (const void*)sym_agg_kernel<8>;
    return (const void*)sym_agg_kernel<16>;
}
static const void* pick_topk_kernel(int world, int use_multimem) {
    if (use_multimem) return (const void*)sym_saliency_topk_kernel<0>;
    if (world <= 2) return (const void*)sym_saliency_topk_kernel<2>;
    if (world <= 4) return (const void*)sym_saliency_topk_kernel<4>;
    if (world <= 8) return (const void*)sym_saliency_topk_kernel<8>;
    return (const void*)sym_saliency_topk_kernel<16>;
}

}  // namespace nidt
using namespace nidt;

// One launch per rank: all-reduce of the score slabs over NVLink, k-th largest normalised score, mask (ties kept).
// scratch: nidt_topk_scratch_bytes() bytes; scores_out (optional) receives the reduced scores.
NIDT_API int nidt_sym_saliency_topk_mask(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial,
                                         void* mc_result, int rank, int world, int64_t n, const unsigned char* flags, float* mask,
                                         unsigned long long k, void* scratch, float* scores_out, uint32_t epoch, int use_multimem, void* stream) {
    NIDT_REQUIRE(n % 4 == 0, "sym_saliency_topk: n must be a multiple of 4");
    NIDT_REQUIRE(scratch != nullptr, "sym_saliency_topk: scratch is null");
    NIDT_REQUIRE(!use_multimem || (mc_partial != nullptr && mc_result != nullptr), "sym_saliency_topk: multimem requested without multicast addresses");
    SymPeers sp;
    if (int rc = fill_peers(sp, partial_ptrs, result_ptrs, flag_ptrs, mc_partial, mc_result, rank, world)) return rc;
    int dev = 0, sms = 0, per_sm = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const void* kern = pick_topk_kernel(world, use_multimem);
    NIDT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kSymThreads, 0));
    NIDT_REQUIRE(per_sm >= 1, "sym_saliency_topk: kernel does not fit on an SM");
    int grid = sms < 2 * kNumSMs ? sms : 2 * kNumSMs;
    TopkScratch* sc = (TopkScratch*)scratch;
    void* args[] = {(void*)&sp, (void*)&n, (void*)&flags, (void*)&mask, (void*)&k, (void*)&sc, (void*)&scores_out, (void*)&epoch};
    NIDT_CUDA_CHECK(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(kSymThreads), args, 0, (cudaStream_t)stream));
    return 0;
}

NIDT_API int nidt_sym_max_ranks() { return kMaxRanks; }
NIDT_API int nidt_sym_flag_words() { return 4 * kMaxRanks; }

// partial_ptrs / result_ptrs / flag_ptrs: HOST arrays of `world` peer-mapped device pointers (same order on all ranks).
// flag pads: nidt_sym_flag_words() uint32 each, zero before the first call; `epoch` must increase by 1 per call
// (identically on every rank).  All ranks must make this call; the kernels wait for each other.
NIDT_API int nidt_sym_reduce_mask_broadcast(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial,
                                            void* mc_result, int rank, int world, const float* mask, int64_t n_mask, int64_t n,
                                            float* out_local, uint32_t epoch, int use_multimem, void* stream) {
    NIDT_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, "sym_reduce: bad rank %d / world %d", rank, world);
    NIDT_REQUIRE(n % 4 == 0 && n_mask % 4 == 0, "sym_reduce: n and n_mask must be multiples of 4");
    NIDT_REQUIRE(!use_multimem || (mc_partial != nullptr && mc_result != nullptr), "sym_reduce: multimem requested without multicast addresses");
    SymPeers sp;
    if (int rc = fill_peers(sp, partial_ptrs, result_ptrs, flag_ptrs, mc_partial, mc_result, rank, world)) return rc;
    int dev = 0, sms = 0, per_sm = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const void* kern = pick_agg_kernel(world, use_multimem);
    NIDT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kSymThreads, 0));
    NIDT_REQUIRE(per_sm >= 1, "sym_reduce: kernel does not fit on an SM");
    int64_t want = (n / 4 / world + 2 * kSymThreads - 1) / (2 * kSymThreads);
    if (want < 1) want = 1;
    const int cap = sms * (per_sm > 2 ? 2 : per_sm);
    int grid = (int)(want < cap ? want : cap);
    void* args[] = {(void*)&sp, (void*)&mask, (void*)&n_mask, (void*)&n, (void*)&out_local, (void*)&epoch};
    NIDT_CUDA_CHECK(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(kSymThreads), args, 0, (cudaStream_t)stream));
    return 0;
}
