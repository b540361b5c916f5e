// Global top-k -> mask (shared by flat_ops.cu and sym_agg.cu).
#pragma once
#include "common.cuh"
#include <cooperative_groups.h>

namespace nidt {
namespace cg = cooperative_groups;

constexpr int kTopkThreads = 512;
struct TopkScratch {           // lives in caller-provided global scratch (zeroed by the kernel)
    unsigned int hist[4][256];
    float block_sums[2 * kNumSMs];
    float threshold;           // out: k-th largest q
    float norm;                // out: sum of prunable scores
    unsigned int kept;         // out: number of prunable entries kept
};

__device__ __forceinline__ void topk_mask_body(
    cg::grid_group& grid, const float* __restrict__ scores, const unsigned char* __restrict__ flags, float* __restrict__ mask,
    int64_t n, unsigned long long k, TopkScratch* __restrict__ sc) {
    __shared__ unsigned int sh_hist[256];
    __shared__ float red[kTopkThreads / 32];
    __shared__ unsigned int sh_prefix;
    __shared__ unsigned long long sh_k;
    const int64_t tid = (int64_t)blockIdx.x * kTopkThreads + threadIdx.x;
    const int64_t nth = (int64_t)gridDim.x * kTopkThreads;

    // phase 0: clear scratch, block-partial sums
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i < 4 * 256; i += kTopkThreads) (&sc->hist[0][0])[i] = 0u;
        if (threadIdx.x == 0) sc->kept = 0u;
    }
    float acc = 0.f;
    for (int64_t i = tid; i < n; i += nth) acc += (flags[i] == 2) ? scores[i] : 0.f;
    float bs = block_sum<kTopkThreads>(acc, red);
    if (threadIdx.x == 0) sc->block_sums[blockIdx.x] = bs;
    grid.sync();

    float p = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += kTopkThreads) p += sc->block_sums[i];
    const float norm = block_sum<kTopkThreads>(p, red);

    unsigned int prefix = 0u;
    unsigned long long krem = k;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        const unsigned int hi_mask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        for (int i = threadIdx.x; i < 256; i += kTopkThreads) sh_hist[i] = 0u;
        __syncthreads();
        for (int64_t i = tid; i < n; i += nth) {
            if (flags[i] == 2) {
                const unsigned int key = __float_as_uint(__fdiv_rn(scores[i], norm));
                if ((key & hi_mask) == prefix) atomicAdd(&sh_hist[(key >> shift) & 0xffu], 1u);
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < 256; i += kTopkThreads)
            if (sh_hist[i]) atomicAdd(&sc->hist[pass][i], sh_hist[i]);
        grid.sync();
        if (threadIdx.x == 0) {
            unsigned long long cum = 0ull;
            int d = 255;
            for (; d > 0; --d) {
                const unsigned long long c = sc->hist[pass][d];
                if (cum + c >= krem) break;
                cum += c;
            }
            sh_prefix = prefix | ((unsigned int)d << shift);
            sh_k = krem - cum;
        }
        __syncthreads();
        prefix = sh_prefix;
        krem = sh_k;
    }
    const float thr = k == 0ull ? __int_as_float(0x7f800000) : __uint_as_float(prefix);

    unsigned int kept = 0u;
    for (int64_t i = tid; i < n; i += nth) {
        const unsigned char f = flags[i];
        float m = 0.f;
        if (f == 1) m = 1.f;
        else if (f == 2) {
            m = (__fdiv_rn(scores[i], norm) >= thr) ? 1.f : 0.f;
            kept += (unsigned int)m;
        }
        mask[i] = m;
    }
    kept = (unsigned int)warp_sum_u64(kept);
    if ((threadIdx.x & 31) == 0 && kept) atomicAdd(&sc->kept, kept);
    if (tid == 0) { sc->threshold = thr; sc->norm = norm; }
}


}  // namespace nidt
