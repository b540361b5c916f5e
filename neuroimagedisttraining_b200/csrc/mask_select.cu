// DisPFL prune / regrow (W10 of SURVEY §8): exact "k smallest |w| among the active weights" / "k largest |g| among the inactive
// ones" on one tensor, as ONE cooperative kernel: 4-pass radix select on the fp32 bit pattern (|x| >= 0, so the bits order like
// the values), then a mask update in which ties at the threshold are taken in index order until exactly k elements changed.
// The reference does a full torch.sort per layer and scatters (fedml_api/standalone/DisPFL/client.py:71-99).
#include "common.cuh"
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace nidt {
namespace {

constexpr int kSelThreads = 512;
constexpr int kSelMaxBlocks = 2 * kNumSMs;

struct SelScratch {
    unsigned int hist[4][256];
    unsigned int block_ties[kSelMaxBlocks];
    unsigned int changed;      // number of mask entries flipped (for the caller / tests)
};

// key: larger = preferred.  mode 0 (fire): candidates mask != 0, prefer SMALL |v|; mode 1 (regrow): candidates mask == 0, prefer LARGE |v|
__device__ __forceinline__ bool candidate(float m, int mode) { return mode == 0 ? (m != 0.f) : (m == 0.f); }
__device__ __forceinline__ unsigned int key_of(float v, int mode) {
    const unsigned int b = __float_as_uint(fabsf(v));
    return mode == 0 ? ~b : b;
}

__global__ void __launch_bounds__(kSelThreads) mask_select_kernel(const float* __restrict__ v, float* __restrict__ mask, int64_t n, unsigned long long k,
                                                                  int mode, SelScratch* __restrict__ sc) {
    cg::grid_group grid = cg::this_grid();
    __shared__ unsigned int sh_hist[256];
    __shared__ unsigned int sh_scan[kSelThreads / 32 + 1];
    __shared__ unsigned int sh_base;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // contiguous chunk per block (index order matters for the ties)
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x, hi = min(n, lo + per);

    if (blockIdx.x == 0) {
        for (int i = tid; i < 4 * 256; i += blockDim.x) (&sc->hist[0][0])[i] = 0u;
        if (tid == 0) sc->changed = 0u;
    }
    grid.sync();

    unsigned int prefix = 0;                 // selected high bytes so far
    unsigned long long remaining = k;        // how many still to take among keys that match the prefix
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        const unsigned int hi_mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
        for (int i = tid; i < 256; i += blockDim.x) sh_hist[i] = 0u;
        __syncthreads();
        for (int64_t i = lo + tid; i < hi; i += blockDim.x) {
            if (!candidate(mask[i], mode)) continue;
            const unsigned int key = key_of(v[i], mode);
            if ((key & hi_mask) == prefix) atomicAdd(&sh_hist[(key >> shift) & 0xFFu], 1u);
        }
        __syncthreads();
        for (int i = tid; i < 256; i += blockDim.x)
            if (sh_hist[i]) atomicAdd(&sc->hist[pass][i], sh_hist[i]);
        grid.sync();
        // every block derives the same digit from the same global histogram
        if (tid == 0) {
            unsigned long long cum = 0;
            int b = 255;
            for (; b > 0; --b) {
                const unsigned long long c = sc->hist[pass][b];
                if (cum + c >= remaining) break;
                cum += c;
            }
            sh_base = (unsigned int)b;
            sh_scan[0] = (unsigned int)(cum & 0xFFFFFFFFull);
            sh_scan[1] = (unsigned int)(cum >> 32);
        }
        __syncthreads();
        prefix |= sh_base << shift;
        remaining -= ((unsigned long long)sh_scan[1] << 32) | sh_scan[0];
        __syncthreads();
    }
    // keys > prefix are all taken; `remaining` of the keys == prefix are taken in index order
    const unsigned int thr = prefix;
    unsigned int my_ties = 0;
    for (int64_t i = lo + tid; i < hi; i += blockDim.x)
        if (candidate(mask[i], mode) && key_of(v[i], mode) == thr) ++my_ties;
    {
        unsigned int w = my_ties;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if (lane == 0) sh_scan[wid] = w;
        __syncthreads();
        if (tid == 0) {
            unsigned int t = 0;
            for (int i = 0; i < kSelThreads / 32; ++i) t += sh_scan[i];
            sc->block_ties[blockIdx.x] = t;
        }
    }
    grid.sync();
    if (tid == 0) {
        unsigned long long before = 0;
        for (unsigned int b = 0; b < blockIdx.x; ++b) before += sc->block_ties[b];
        sh_scan[0] = (unsigned int)(before & 0xFFFFFFFFull);
        sh_scan[1] = (unsigned int)(before >> 32);
    }
    __syncthreads();
    unsigned long long tie_rank = ((unsigned long long)sh_scan[1] << 32) | sh_scan[0];     // ties in earlier blocks
    __syncthreads();
    const float new_val = mode == 0 ? 0.f : 1.f;
    unsigned int flipped = 0;
    for (int64_t base = lo; base < hi; base += blockDim.x) {
        const int64_t i = base + tid;
        bool cand = false, tie = false, take = false;
        if (i < hi && candidate(mask[i], mode)) {
            cand = true;
            const unsigned int key = key_of(v[i], mode);
            tie = key == thr;
            take = key > thr;
        }
        // ordered rank of this thread's tie inside the block-wide batch
        const unsigned int bal = __ballot_sync(0xffffffffu, tie);
        const unsigned int in_warp = __popc(bal & ((1u << lane) - 1u));
        if (lane == 0) sh_scan[wid] = __popc(bal);
        __syncthreads();
        unsigned int before_warp = 0, batch_total = 0;
        for (int w = 0; w < kSelThreads / 32; ++w) {
            const unsigned int c = sh_scan[w];
            if (w < wid) before_warp += c;
            batch_total += c;
        }
        if (tie && tie_rank + before_warp + in_warp < remaining) take = true;
        if (cand && take) { mask[i] = new_val; ++flipped; }
        tie_rank += batch_total;
        __syncthreads();
    }
    if (flipped) atomicAdd(&sc->changed, flipped);
}

}  // namespace
}  // namespace nidt
using namespace nidt;

NIDT_API int nidt_mask_select_scratch_bytes() { return (int)sizeof(SelScratch); }

// mask (fp32 0/1, updated in place): mode 0 clears the k active entries with the smallest |v|, mode 1 sets the k inactive entries with
// the largest |v|; ties at the threshold go in index order, so exactly min(k, #candidates) entries change.  scratch:
// nidt_mask_select_scratch_bytes() bytes; after the call its last word holds the number of entries changed
// (nidt_mask_select_read_changed).  ref: DisPFL/client.py:71-99 (torch.sort per layer).
NIDT_API int nidt_mask_select_k(const float* v, float* mask, int64_t n, int64_t k, int mode, void* scratch, void* stream) {
    NIDT_REQUIRE(n >= 0 && k >= 0 && (mode == 0 || mode == 1) && scratch != nullptr, "nidt_mask_select_k: bad arguments (n=%lld k=%lld mode=%d)",
                 (long long)n, (long long)k, mode);
    if (n == 0 || k == 0) {
        NIDT_CUDA_CHECK(cudaMemsetAsync(&((SelScratch*)scratch)->changed, 0, sizeof(unsigned int), (cudaStream_t)stream));
        return 0;
    }
    int dev = 0, sms = 0, per_sm = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    NIDT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mask_select_kernel, kSelThreads, 0));
    NIDT_REQUIRE(per_sm >= 1, "nidt_mask_select_k: kernel does not fit on an SM");
    int64_t want = (n + 4 * kSelThreads - 1) / (4 * kSelThreads);
    int grid = sms * (per_sm > 2 ? 2 : per_sm);
    if (grid > kSelMaxBlocks) grid = kSelMaxBlocks;
    if (want < grid) grid = (int)(want < 1 ? 1 : want);
    SelScratch* sc = (SelScratch*)scratch;
    unsigned long long kk = (unsigned long long)k;
    void* args[] = {(void*)&v, (void*)&mask, (void*)&n, (void*)&kk, (void*)&mode, (void*)&sc};
    NIDT_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)mask_select_kernel, dim3(grid), dim3(kSelThreads), args, 0, (cudaStream_t)stream));
    return 0;
}

// number of mask entries the last nidt_mask_select_k on this scratch changed (synchronises the stream)
NIDT_API int nidt_mask_select_read_changed(const void* scratch, unsigned int* changed, void* stream) {
    NIDT_CUDA_CHECK(cudaMemcpyAsync(changed, &((const SelScratch*)scratch)->changed, sizeof(unsigned int), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    NIDT_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}
