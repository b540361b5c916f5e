// Cross-GPU aggregation fused with its collective (north-star kernels W1 / W2 of SURVEY §8).
//
// W1 (sym_agg_kernel): ONE cooperative kernel per rank does
//     local weighting : partial = sum_c (n_c / sum n) * w_c over this rank's K sampled clients, multiply and add kept apart in
//                       client order (bit-identical to the reference's loop), written to NVLink-mapped symmetric memory;
//     reduce-scatter  : rank r sums chunk r of all ranks' partials -- P2P loads in fixed rank order (every rank computes
//                       bit-identical values), the same through shared memory with bulk-async copies from the peers
//                       (cp.async.bulk, no register cost per byte in flight), or multimem.ld_reduce (NVLS, in-switch reduction);
//     mask            : multiplies the SalientGrads mask into the reduced chunk;
//     all-gather      : writes the chunk into every rank's `result` buffer (P2P stores, bulk stores from shared memory, or one
//                       multimem.st): the new global model lands in all slabs at once.  When the caller's global slab IS the
//                       result buffer (SymmetricAggregator.adopt) nothing else is left to do; otherwise a write-back copies it.
//   A chunk can also be split between the in-switch path and a P2P path (`mm_frac`): the NVLS reduction unit and the plain
//   load/store path are different resources of the same links.
// Cross-rank ordering uses release/acquire flags in symmetric memory with a monotonically increasing epoch (no resets),
// grid-wide ordering uses a cooperative launch.  No NCCL call on the data path.
// Reference: the CPU dict loop `_aggregate` (fedml_api/standalone/sailentgrads/sailentgrads_api.py:212-227).
//
// W2 (sym_saliency_topk_kernel): the score slabs are reduce-scattered the same way; every rank then keeps ONLY its shard:
// it histograms the shard's fp32 bit patterns (3 radix passes of 11/11/10 bits), the 2048-bin histograms (and the shard sums)
// are exchanged through symmetric memory and added in rank order, so all ranks agree on the k-th largest score without anyone
// holding all scores; each rank emits the mask of its shard into every rank's mask buffer (all-gather by P2P stores).
// `replicate`: for slabs that fit in L2 the three dependent histogram exchanges (one cross-rank barrier each, ~10 us at 8 GPUs) cost more
// than the passes themselves; the kernel then all-gathers the reduced scores (as W1 does) and every rank selects over the whole slab on
// its own -- 2 cross-rank barriers per call instead of 6, same kernel, same buffers.
// Reference: snip.py:80-140 (mean over clients, cat, /sum, topk, >= threshold).
#include "common.cuh"
#include "tc_common.cuh"
#include "topk.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>

namespace nidt {

constexpr int kMaxRanks = 16;
constexpr int kSymThreads = 512;
constexpr int kMaxLocal = 48;          // client slabs weighted inside the kernel (more: the caller pre-reduces)
constexpr int kRadixBits0 = 11, kRadixBits2 = 10;
constexpr int kRadixBins = 2048;
constexpr int kHistWords = 3 * kMaxRanks * kRadixBins + 2 * kMaxRanks;   // [pass][src rank][bin] counts + per-rank shard sums (float bits)

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

struct AggLocal {                   // this rank's clients, weighted inside the kernel (k == 0: `partial` is already complete)
    const float* src[kMaxLocal];
    float wt[kMaxLocal];
    int k;
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
    if (order_prior_writes) {      // the data this rank publishes was written by this grid
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
// shared -> global bulk copy (the destination may be a peer's memory), tracked by the bulk async-group of the issuing thread
__device__ __forceinline__ void bulk_store_1d(void* dst, const void* src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(tc::smem_u32(src_smem)), "r"(bytes) : "memory");
}

// ------------------------------------------------------------------------------------------ phase A: local weighting
// partial[i] = sum_c wt[c] * src[c][i] over the whole slab, multiply and add separate in client order (flat_ops.cu does the same)
__device__ __forceinline__ void local_weighted_sum(const AggLocal& a, float* __restrict__ partial, int64_t n) {
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 v = ld4(a.src[0] + 4 * i);
        float wt = a.wt[0];
        float4 acc;
        acc.x = __fmul_rn(v.x, wt); acc.y = __fmul_rn(v.y, wt); acc.z = __fmul_rn(v.z, wt); acc.w = __fmul_rn(v.w, wt);
#pragma unroll 4
        for (int s = 1; s < a.k; ++s) {
            v = ld4(a.src[s] + 4 * i);
            wt = a.wt[s];
            acc.x = __fadd_rn(acc.x, __fmul_rn(v.x, wt)); acc.y = __fadd_rn(acc.y, __fmul_rn(v.y, wt));
            acc.z = __fadd_rn(acc.z, __fmul_rn(v.z, wt)); acc.w = __fadd_rn(acc.w, __fmul_rn(v.w, wt));
        }
        st4(partial + 4 * i, acc);
    }
}

// ------------------------------------------------------------------------------------------ phase B variants
// Every variant reduces float4 range [lo, hi) of the slab over all partials (+mask) and writes it into every result buffer
// (`broadcast`) or only into this rank's (W2 keeps the shard), using CTAs [cta0, cta0 + nctas) of the grid.

// P2P through registers.  kWorld (2/4/8/16 >= world) bounds the registers; at <= 4 ranks two vectors per peer are in flight.
template <int kWorld>
__device__ __forceinline__ void rs_ag_p2p(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t lo, int64_t hi, int cta0, int nctas,
                                          bool broadcast) {
    const int64_t tid = (int64_t)((int)blockIdx.x - cta0) * blockDim.x + threadIdx.x, nth = (int64_t)nctas * blockDim.x;
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
        if (broadcast) {
#pragma unroll
            for (int p = 0; p < kWorld; ++p)
                if (p < sp.world) {
                    st4(sp.result[p] + 4 * i0, s0);
                    if (two) st4(sp.result[p] + 4 * i1, s1);
                }
        } else {
            st4(sp.result[sp.rank] + 4 * i0, s0);
            if (two) st4(sp.result[sp.rank] + 4 * i1, s1);
        }
    }
}

// NVLS: in-switch reduction on the pull, one multicast store on the push; U independent 16-byte requests per thread in flight
template <int U>
__device__ __forceinline__ void rs_ag_multimem(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t lo, int64_t hi, int cta0,
                                               int nctas, bool broadcast) {
    // a CTA works on contiguous tiles of U * blockDim float4 (32 KiB at U = 4), tiles round-robin over the CTAs: the requests a
    // CTA has in flight stay within a few pages on every GPU the switch reads from
    const int64_t tile = (int64_t)U * blockDim.x, nth = blockDim.x;
    for (int64_t i = lo + (int64_t)((int)blockIdx.x - cta0) * tile + threadIdx.x; i < hi; i += (int64_t)nctas * tile) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (i + u * nth < hi) v[u] = multimem_ld_reduce_f32x4(sp.mc_partial + 4 * (i + u * nth));
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t j = i + u * nth;
            if (j < hi) {
                if (mask != nullptr && 4 * j < n_mask) { const float4 mk = ld4(mask + 4 * j); v[u].x *= mk.x; v[u].y *= mk.y; v[u].z *= mk.z; v[u].w *= mk.w; }
                if (broadcast) multimem_st_f32x4(sp.mc_result + 4 * j, v[u]);
                else st4(sp.result[sp.rank] + 4 * j, v[u]);
            }
        }
    }
}

// P2P through shared memory: one thread queues bulk-async copies of a tile from every peer (kBulkStages tiles in flight per CTA,
// no registers tied up by bytes in flight), all threads add the P copies in rank order, the masked tile goes back out with one
// bulk store per peer.  Dynamic shared memory: kBulkStages * world * kBulkTileBytes + 3 * kBulkTileBytes + barriers.
constexpr int kBulkTileF4 = kSymThreads;                  // one float4 per thread and tile
constexpr int kBulkTileBytes = kBulkTileF4 * 16;          // 8 KiB
constexpr int kBulkStages = 3;
__host__ __device__ constexpr size_t bulk_smem_bytes(int world) { return (size_t)kBulkStages * world * kBulkTileBytes + 3 * (size_t)kBulkTileBytes + 64; }

__device__ __forceinline__ void rs_ag_bulk(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t lo, int64_t hi, int cta0, int nctas,
                                           bool broadcast, unsigned char* smem) {
    const int P = sp.world;
    float4* stage = reinterpret_cast<float4*>(smem);                                            // [kBulkStages][P][kBulkTileF4]
    float4* outb = stage + (size_t)kBulkStages * P * kBulkTileF4;                               // [3][kBulkTileF4]
    uint64_t* full = reinterpret_cast<uint64_t*>(outb + 3 * kBulkTileF4);                       // [kBulkStages]
    const int cta = (int)blockIdx.x - cta0;
    const int64_t n_tiles = hi > lo ? (hi - lo + kBulkTileF4 - 1) / kBulkTileF4 : 0;
    const int64_t my_tiles = cta < n_tiles ? (n_tiles - cta + nctas - 1) / nctas : 0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kBulkStages; ++s) tc::mbar_init(full + s, 1);
        tc::mbar_fence_init();
    }
    __syncthreads();
    auto issue = [&](int64_t j) {                       // thread 0: queue the loads of my j-th tile
        const int64_t t0 = lo + ((int64_t)cta + j * nctas) * kBulkTileF4;
        const int64_t cnt = min((int64_t)kBulkTileF4, hi - t0);
        const int s = (int)(j % kBulkStages);
        tc::mbar_expect_tx(full + s, (uint32_t)(cnt * 16 * P));
        for (int p = 0; p < P; ++p)
            tc::bulk_load_1d(stage + ((size_t)s * P + p) * kBulkTileF4, sp.partial[p] + 4 * t0, (uint32_t)(cnt * 16), full + s);
    };
    if (threadIdx.x == 0)
        for (int64_t j = 0; j < kBulkStages - 1 && j < my_tiles; ++j) issue(j);
    for (int64_t j = 0; j < my_tiles; ++j) {
        const int s = (int)(j % kBulkStages);
        if (threadIdx.x == 0) {
            if (j + kBulkStages - 1 < my_tiles) issue(j + kBulkStages - 1);     // its stage was consumed in iteration j-1 (barrier below)
            tc::tma_store_wait_read_n<1>();             // at most the previous tile's store still reads shared memory
        }
        tc::mbar_wait(full + s, (uint32_t)((j / kBulkStages) & 1));
        const int64_t t0 = lo + ((int64_t)cta + j * nctas) * kBulkTileF4;
        const int64_t cnt = min((int64_t)kBulkTileF4, hi - t0);
        float4* ob = outb + (size_t)(j % 3) * kBulkTileF4;                      // last read by the store of tile j-3: drained (see above)
        if (threadIdx.x < cnt) {
            const float4* st = stage + (size_t)s * P * kBulkTileF4 + threadIdx.x;
            float4 acc = st[0];
            for (int p = 1; p < P; ++p) {
                const float4 v = st[(size_t)p * kBulkTileF4];
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            const int64_t i = t0 + threadIdx.x;
            if (mask != nullptr && 4 * i < n_mask) { const float4 mk = ld4(mask + 4 * i); acc.x *= mk.x; acc.y *= mk.y; acc.z *= mk.z; acc.w *= mk.w; }
            ob[threadIdx.x] = acc;
        }
        tc::fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) {
            if (broadcast) for (int p = 0; p < P; ++p) bulk_store_1d(sp.result[p] + 4 * t0, ob, (uint32_t)(cnt * 16));
            else bulk_store_1d(sp.result[sp.rank] + 4 * t0, ob, (uint32_t)(cnt * 16));
            tc::tma_store_commit();
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // this CTA's stores are complete (written, not just read)
    __syncthreads();
}

enum SymMode { kModeP2P = 0, kModeMultimem = 1, kModeBulk = 2 };

// Chunk `rank` of the slab, optionally split: the first mm_frac_q8/256 of it through the in-switch path on a matching share
// of the CTAs, the rest through the P2P path `mode`.
template <int kWorld>
__device__ __forceinline__ void reduce_mask_broadcast(const SymPeers& sp, const float* __restrict__ mask, int64_t n_mask, int64_t n, int mode,
                                                      int mm_frac_q8, bool broadcast, unsigned char* smem) {
    const int64_t n4 = n >> 2;
    const int64_t per = (n4 + sp.world - 1) / sp.world;
    const int64_t lo = per * sp.rank, hi = min(n4, lo + per);
    const int G = (int)gridDim.x;
    if (mode == kModeMultimem || mm_frac_q8 >= 256) { rs_ag_multimem<8>(sp, mask, n_mask, lo, hi, 0, G, broadcast); return; }
    int g_mm = 0;
    int64_t mid = lo;
    if (mm_frac_q8 > 0 && sp.mc_partial != nullptr && G >= 2 && hi > lo) {
        g_mm = max(1, min(G - 1, (G * mm_frac_q8) >> 8));
        mid = lo + (((hi - lo) * mm_frac_q8) >> 8);
    }
    if ((int)blockIdx.x < g_mm) { rs_ag_multimem<8>(sp, mask, n_mask, lo, mid, 0, g_mm, broadcast); return; }
    if (mode == kModeBulk) rs_ag_bulk(sp, mask, n_mask, mid, hi, g_mm, G - g_mm, broadcast, smem);
    else rs_ag_p2p<kWorld>(sp, mask, n_mask, mid, hi, g_mm, G - g_mm, broadcast);
}

template <int kWorld>
__global__ void __launch_bounds__(kSymThreads, kWorld <= 8 ? 2 : 1) sym_agg_kernel(const SymPeers sp, const __grid_constant__ AggLocal local,
                                                                                   const float* __restrict__ mask, int64_t n_mask, int64_t n,
                                                                                   float* __restrict__ out_local, uint32_t epoch, int mode, int mm_frac_q8) {
    extern __shared__ __align__(128) unsigned char sym_smem[];
    cg::grid_group grid = cg::this_grid();
    if (local.k > 0) local_weighted_sum(local, sp.partial[sp.rank], n);
    cross_rank_barrier(grid, sp, 0, 2 * epoch + 1, local.k > 0);   // every rank's partial is complete
    reduce_mask_broadcast<kWorld>(sp, mask, n_mask, n, mode, mm_frac_q8, true, sym_smem);
    cross_rank_barrier(grid, sp, 0, 2 * epoch + 2);                 // every chunk has landed everywhere
    if (out_local != nullptr && out_local != sp.result[sp.rank]) {
        const float* src = sp.result[sp.rank];
        const int64_t n4 = n >> 2;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
            st4(out_local + 4 * i, ld4(src + 4 * i));
    }
}

// ------------------------------------------------------------------------------------------ W2
struct SalAux {
    uint32_t* hist[kMaxRanks];      // peer-mapped exchange pads: [pass][src rank][bin] counts, then the shard sums
    float* mask_out[kMaxRanks];     // peer-mapped mask buffers (symmetric), all-gathered by P2P stores
    uint32_t* local_hist;           // this rank's scratch: 3 * kRadixBins counters + 1 kept counter (zeroed by the kernel)
    float* block_sums;              // 2*kNumSMs floats
};
struct SalResult { float threshold, norm; unsigned int kept; unsigned int pad; };

__device__ __forceinline__ int radix_shift(int pass) { return pass == 0 ? 21 : (pass == 1 ? 10 : 0); }
__device__ __forceinline__ int radix_bits(int pass) { return pass == 2 ? kRadixBits2 : kRadixBits0; }

// `partial` holds this rank's sum of its clients' (already 1/num_clients scaled) |g*w| scores; flags: 0 padding, 1 never pruned, 2 prunable.
template <int kWorld>
__global__ void __launch_bounds__(kSymThreads) sym_saliency_topk_kernel(const SymPeers sp, const SalAux aux, int64_t n, const unsigned char* __restrict__ flags,
                                                                        float* __restrict__ mask, unsigned long long k, SalResult* __restrict__ res,
                                                                        float* __restrict__ scores_out, uint32_t epoch, int mode, int replicate) {
    extern __shared__ __align__(128) unsigned char sym_smem[];
    __shared__ unsigned int sh_hist[kRadixBins];
    __shared__ unsigned long long sh_scan[kSymThreads / 32];
    __shared__ float red[kSymThreads / 32];
    __shared__ unsigned int sh_prefix;
    __shared__ unsigned long long sh_k;
    cg::grid_group grid = cg::this_grid();
    const int P = sp.world;
    const int64_t n4 = n >> 2;
    const int64_t per4 = (n4 + P - 1) / P;
    // the range this rank selects over: its shard, or everything (replicate)
    const int64_t lo = replicate ? 0 : min(n, 4 * per4 * sp.rank), hi = replicate ? n : min(n, lo + 4 * per4);
    const int Ps = replicate ? 1 : P;                                                 // histograms added per pass
    const int64_t tid = (int64_t)blockIdx.x * kSymThreads + threadIdx.x, nth = (int64_t)gridDim.x * kSymThreads;

    // clear this call's scratch, then reduce-scatter: the reduced shard stays HERE (result[rank][lo:hi))
    for (int64_t i = tid; i < 3 * kRadixBins + 1; i += nth) aux.local_hist[i] = 0u;
    uint32_t bar = 6 * epoch;
    cross_rank_barrier(grid, sp, 0, ++bar, false);                            // every rank's partial is complete
    reduce_mask_broadcast<kWorld>(sp, nullptr, 0, n, mode, 0, replicate != 0, sym_smem);
    if (replicate) {
        cross_rank_barrier(grid, sp, 0, ++bar);                               // every chunk of reduced scores has landed here
    } else {
        __threadfence();
        grid.sync();
    }
    const float* scores = sp.result[sp.rank];

    // shard sum of the prunable scores (fixed-order block partials)
    float acc = 0.f;
    for (int64_t i = lo + tid; i < hi; i += nth) acc += (flags[i] == 2) ? scores[i] : 0.f;
    const float bs = block_sum<kSymThreads>(acc, red);
    if (threadIdx.x == 0) aux.block_sums[blockIdx.x] = bs;

    unsigned int prefix = 0u;
    unsigned long long krem = k;
    float norm = 0.f;
    for (int pass = 0; pass < 3; ++pass) {
        const int shift = radix_shift(pass), bits = radix_bits(pass);
        const unsigned int hi_mask = pass == 0 ? 0u : (0xffffffffu << (shift + bits));
        for (int i = threadIdx.x; i < kRadixBins; i += kSymThreads) sh_hist[i] = 0u;
        if (threadIdx.x == 0) { sh_prefix = prefix; sh_k = krem; }
        __syncthreads();
        for (int64_t i = lo + tid; i < hi; i += nth)
            if (flags[i] == 2) {
                const unsigned int key = __float_as_uint(scores[i]);            // scores >= 0: the bit pattern is order preserving
                if ((key & hi_mask) == prefix) atomicAdd(&sh_hist[(key >> shift) & ((1u << bits) - 1u)], 1u);
            }
        __syncthreads();
        uint32_t* lh = aux.local_hist + pass * kRadixBins;
        for (int i = threadIdx.x; i < kRadixBins; i += kSymThreads)
            if (sh_hist[i]) atomicAdd(lh + i, sh_hist[i]);
        __threadfence();
        grid.sync();
        // publish the shard histogram (and, with pass 0, the shard sum) into every rank's exchange pad
        if (replicate) {
            if (pass == 0) {                                                   // everybody adds the block sums in the same order
                float p = 0.f;
                for (int i = threadIdx.x; i < (int)gridDim.x; i += kSymThreads) p += aux.block_sums[i];
                norm = block_sum<kSymThreads>(p, red);
            }
        } else {
        if (pass == 0 && blockIdx.x == 0) {
            float p = 0.f;
            for (int i = threadIdx.x; i < (int)gridDim.x; i += kSymThreads) p += aux.block_sums[i];
            const float total = block_sum<kSymThreads>(p, red);
            if (threadIdx.x < P) aux.hist[threadIdx.x][3 * kMaxRanks * kRadixBins + sp.rank] = __float_as_uint(total);
        }
        for (int64_t i = tid; i < (int64_t)P * kRadixBins; i += nth) {
            const int dst = (int)(i / kRadixBins), b = (int)(i % kRadixBins);
            aux.hist[dst][((size_t)pass * kMaxRanks + sp.rank) * kRadixBins + b] = lh[b];
        }
        cross_rank_barrier(grid, sp, 0, ++bar);
        }
        // every block finds the digit itself from the published histograms (no extra grid-wide step)
        const uint32_t* hp = replicate ? lh : aux.hist[sp.rank] + (size_t)pass * kMaxRanks * kRadixBins;
        constexpr int kPer = kRadixBins / kSymThreads;                         // 4 consecutive bins per thread
        unsigned int c[kPer];
        unsigned long long mine = 0ull;
#pragma unroll
        for (int q = 0; q < kPer; ++q) {
            unsigned int s = 0u;
            for (int r = 0; r < Ps; ++r) s += hp[(size_t)r * kRadixBins + threadIdx.x * kPer + q];
            c[q] = s;
            mine += s;
        }
        // suffix sums over threads: above = number of keys in the bins held by threads > me
        unsigned long long incl = mine;
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long v = __shfl_down_sync(0xffffffffu, incl, o);
            if (lane + o < 32) incl += v;
        }
        if (lane == 0) sh_scan[wid] = incl;
        __syncthreads();
        unsigned long long above = incl - mine;
        for (int w = wid + 1; w < kSymThreads / 32; ++w) above += sh_scan[w];
        if (krem > 0ull && above < krem && above + mine >= krem) {             // the k-th largest key of this pass lies in my bins
            unsigned long long cum = above;
            int q = kPer - 1;
            for (; q > 0; --q) {
                if (cum + c[q] >= krem) break;
                cum += c[q];
            }
            sh_prefix = prefix | ((unsigned int)(threadIdx.x * kPer + q) << shift);
            sh_k = krem - cum;
        }
        __syncthreads();
        prefix = sh_prefix;
        krem = sh_k;
        if (pass == 0 && !replicate) {                                         // shard sums arrived with the first exchange; add in rank order
            const uint32_t* sums = aux.hist[sp.rank] + 3 * kMaxRanks * kRadixBins;
            for (int r = 0; r < P; ++r) norm += __uint_as_float(sums[r]);
        }
        __syncthreads();
    }
    const float thr_q = k == 0ull ? __int_as_float(0x7f800000) : __fdiv_rn(__uint_as_float(prefix), norm);

    // mask of my shard -> every rank's mask buffer
    unsigned int kept = 0u;
    for (int64_t i4 = (lo >> 2) + tid; i4 < (hi >> 2); i4 += nth) {             // shard bounds are multiples of 4: 16-byte stores to every peer
        const uchar4 f4 = *reinterpret_cast<const uchar4*>(flags + 4 * i4);
        const float4 s4 = ld4(scores + 4 * i4);
        const unsigned char f[4] = {f4.x, f4.y, f4.z, f4.w};
        const float sv[4] = {s4.x, s4.y, s4.z, s4.w};
        float m[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            m[q] = 0.f;
            if (f[q] == 1) m[q] = 1.f;
            else if (f[q] == 2) {
                m[q] = (__fdiv_rn(sv[q], norm) >= thr_q) ? 1.f : 0.f;
                kept += (unsigned int)m[q];
            }
        }
        const float4 mv = make_float4(m[0], m[1], m[2], m[3]);
        if (replicate) st4(aux.mask_out[sp.rank] + 4 * i4, mv);
        else for (int p = 0; p < P; ++p) st4(aux.mask_out[p] + 4 * i4, mv);
    }
    kept = (unsigned int)warp_sum_u64(kept);
    if ((threadIdx.x & 31) == 0 && kept) atomicAdd(aux.local_hist + 3 * kRadixBins, kept);
    if (scores_out != nullptr && !replicate) {            // all-gather of the reduced scores through the result buffers (api.scores / diagnostics)
        for (int64_t i = (lo >> 2) + tid; i < (hi >> 2); i += nth) {
            const float4 v = ld4(scores + 4 * i);
            for (int p = 0; p < P; ++p)
                if (p != sp.rank) st4(sp.result[p] + 4 * i, v);
        }
    }
    // the kept counts travel through the (now free) pass-1 rows of the exchange pads
    __threadfence();
    grid.sync();
    if (!replicate) {
        if (blockIdx.x == 0 && threadIdx.x < P) aux.hist[threadIdx.x][((size_t)1 * kMaxRanks + sp.rank) * kRadixBins] = aux.local_hist[3 * kRadixBins];
        cross_rank_barrier(grid, sp, 0, ++bar);
    }
    if (mask != aux.mask_out[sp.rank])
        for (int64_t i = tid; i < n4; i += nth) st4(mask + 4 * i, ld4(aux.mask_out[sp.rank] + 4 * i));
    if (scores_out != nullptr && scores_out != scores)
        for (int64_t i = tid; i < n4; i += nth) st4(scores_out + 4 * i, ld4(scores + 4 * i));
    if (tid == 0) {
        unsigned int total = 0u;
        if (replicate) total = aux.local_hist[3 * kRadixBins];
        else for (int r = 0; r < P; ++r) total += aux.hist[sp.rank][((size_t)1 * kMaxRanks + r) * kRadixBins];
        res->threshold = thr_q; res->norm = norm; res->kept = total;
    }
    // one more line-up so that nobody starts the next call's pad writes while a slower rank still reads this call's.  (replicate: nothing
    // of this rank is written by a peer after the all-gather barrier, and the next call's writes into it come after ITS first barrier,
    // which needs this rank to have left this call.)
    if (!replicate) cross_rank_barrier(grid, sp, 0, ++bar);
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

static const void* pick_agg_kernel(int world) {
    if (world <= 2) return (const void*)sym_agg_kernel<2>;
    if (world <= 4) return (const void*)sym_agg_kernel<4>;
    if (world <= 8) return (const void*)sym_agg_kernel<8>;
    return (const void*)sym_agg_kernel<16>;
}
static const void* pick_topk_kernel(int world) {
    if (world <= 2) return (const void*)sym_saliency_topk_kernel<2>;
    if (world <= 4) return (const void*)sym_saliency_topk_kernel<4>;
    if (world <= 8) return (const void*)sym_saliency_topk_kernel<8>;
    return (const void*)sym_saliency_topk_kernel<16>;
}

}  // namespace nidt
using namespace nidt;

NIDT_API int nidt_sym_max_ranks() { return kMaxRanks; }
NIDT_API int nidt_sym_flag_words() { return 4 * kMaxRanks; }
NIDT_API int nidt_sym_max_local() { return kMaxLocal; }
NIDT_API int nidt_sym_hist_words() { return kHistWords; }
NIDT_API int nidt_sym_bulk_max_ranks() { return (int)((227 * 1024 - 64 - 3 * kBulkTileBytes) / (kBulkStages * kBulkTileBytes)); }
NIDT_API int nidt_sym_sal_scratch_bytes() { return (int)(64 + (3 * kRadixBins + 4) * sizeof(uint32_t) + 2 * kNumSMs * sizeof(float) + 64); }

// W1.  partial_ptrs / result_ptrs / flag_ptrs: HOST arrays of `world` peer-mapped device pointers (same order on all ranks).
// flag pads: nidt_sym_flag_words() uint32 each, zero before the first call; `epoch` must increase by 1 per call (identically on
// every rank).  local_srcs / local_wts: this rank's k_local client slabs and weights (k_local == 0: partial_ptrs[rank] already
// holds the rank's contribution).  mode: 0 P2P through registers, 1 NVLS multimem, 2 P2P through shared memory (bulk-async);
// mm_frac_q8 (0..256): share of every chunk that goes through the in-switch path next to the P2P path; ctas: grid size, 0 = auto.
// out_local == NULL or == result_ptrs[rank]: the result stays in the symmetric buffer (no write-back pass).
// All ranks must make this call; the kernels wait for each other.
NIDT_API int nidt_sym_aggregate(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial, void* mc_result,
                                int rank, int world, const float* const* local_srcs, const float* local_wts, int k_local, const float* mask,
                                int64_t n_mask, int64_t n, float* out_local, uint32_t epoch, int mode, int mm_frac_q8, int ctas, void* stream) {
    NIDT_REQUIRE(n % 4 == 0 && n_mask % 4 == 0, "sym_aggregate: n and n_mask must be multiples of 4");
    NIDT_REQUIRE(k_local >= 0 && k_local <= kMaxLocal, "sym_aggregate: %d local sources (max %d: pre-reduce the rest)", k_local, kMaxLocal);
    NIDT_REQUIRE(mode >= 0 && mode <= 2 && mm_frac_q8 >= 0 && mm_frac_q8 <= 256, "sym_aggregate: bad mode %d / mm_frac %d", mode, mm_frac_q8);
    NIDT_REQUIRE(!(mode == kModeMultimem || mm_frac_q8 > 0) || (mc_partial != nullptr && mc_result != nullptr),
                 "sym_aggregate: multimem requested without multicast addresses");
    SymPeers sp;
    if (int rc = fill_peers(sp, partial_ptrs, result_ptrs, flag_ptrs, mc_partial, mc_result, rank, world)) return rc;
    AggLocal loc;
    loc.k = k_local;
    for (int i = 0; i < kMaxLocal; ++i) { loc.src[i] = i < k_local ? local_srcs[i] : nullptr; loc.wt[i] = i < k_local ? local_wts[i] : 0.f; }
    int dev = 0, sms = 0, per_sm = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const void* kern = pick_agg_kernel(world);
    const size_t smem = mode == kModeBulk ? bulk_smem_bytes(world) : 0;
    NIDT_REQUIRE(smem <= 227 * 1024, "sym_aggregate: the shared-memory path supports up to %d ranks", nidt_sym_bulk_max_ranks());
    if (smem > 0) NIDT_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    NIDT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kSymThreads, smem));
    NIDT_REQUIRE(per_sm >= 1, "sym_aggregate: kernel does not fit on an SM");
    const int cap = sms * (per_sm > 2 ? 2 : per_sm);
    int grid;
    if (ctas > 0) grid = ctas < cap ? ctas : cap;
    else {
        // enough CTAs to cover the larger of the two phases with ~2 float4 per thread, capped at what is co-resident
        const int64_t work4 = k_local > 0 ? n / 4 : n / 4 / world;
        int64_t want = (work4 + 2 * kSymThreads - 1) / (2 * kSymThreads);
        if (want < 1) want = 1;
        grid = (int)(want < cap ? want : cap);
    }
    if (mm_frac_q8 > 0 && mm_frac_q8 < 256 && grid < 2 && cap >= 2) grid = 2;
    void* args[] = {(void*)&sp, (void*)&loc, (void*)&mask, (void*)&n_mask, (void*)&n, (void*)&out_local, (void*)&epoch, (void*)&mode, (void*)&mm_frac_q8};
    NIDT_CUDA_CHECK(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(kSymThreads), args, smem, (cudaStream_t)stream));
    return 0;
}

// the first interface of W1 (partial already in place, P2P registers or multimem), kept for callers of the C ABI
NIDT_API int nidt_sym_reduce_mask_broadcast(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial,
                                            void* mc_result, int rank, int world, const float* mask, int64_t n_mask, int64_t n,
                                            float* out_local, uint32_t epoch, int use_multimem, void* stream) {
    return nidt_sym_aggregate(partial_ptrs, result_ptrs, flag_ptrs, mc_partial, mc_result, rank, world, nullptr, nullptr, 0, mask, n_mask, n, out_local,
                              epoch, use_multimem ? kModeMultimem : kModeP2P, 0, 0, stream);
}

// W2.  One launch per rank: reduce-scatter of the score slabs over NVLink, global k-th largest score from exchanged shard
// histograms, mask all-gather (ties kept).  hist_ptrs: peer-mapped exchange pads of nidt_sym_hist_words() uint32;
// mask_ptrs: peer-mapped mask buffers of n floats; scratch: nidt_sym_sal_scratch_bytes() bytes of local memory (the result
// {threshold, norm, kept} is left at its start for nidt_sym_saliency_read_result); scores_out (optional) receives the reduced
// scores.  `epoch` must increase by 1 per call on the flag pad used here (a pad of its own, not W1's: 6 barriers per call).
NIDT_API int nidt_sym_saliency_topk_mask(void* const* partial_ptrs, void* const* result_ptrs, void* const* flag_ptrs, void* mc_partial,
                                         void* mc_result, void* const* hist_ptrs, void* const* mask_ptrs, int rank, int world, int64_t n,
                                         const unsigned char* flags, float* mask, unsigned long long k, void* scratch, float* scores_out,
                                         uint32_t epoch, int use_multimem, void* stream) {
    NIDT_REQUIRE(n % 4 == 0, "sym_saliency_topk: n must be a multiple of 4");
    NIDT_REQUIRE(scratch != nullptr && hist_ptrs != nullptr && mask_ptrs != nullptr, "sym_saliency_topk: scratch / exchange pads missing");
    NIDT_REQUIRE(!use_multimem || mc_partial != nullptr, "sym_saliency_topk: multimem requested without multicast addresses");
    SymPeers sp;
    if (int rc = fill_peers(sp, partial_ptrs, result_ptrs, flag_ptrs, mc_partial, mc_result, rank, world)) return rc;
    SalAux aux;
    for (int p = 0; p < kMaxRanks; ++p) {
        aux.hist[p] = p < world ? (uint32_t*)hist_ptrs[p] : nullptr;
        aux.mask_out[p] = p < world ? (float*)mask_ptrs[p] : nullptr;
    }
    unsigned char* sc = (unsigned char*)scratch;
    SalResult* res = (SalResult*)sc;
    aux.local_hist = (uint32_t*)(sc + 64);
    aux.block_sums = (float*)(sc + 64 + (3 * kRadixBins + 4) * sizeof(uint32_t));
    int dev = 0, sms = 0, per_sm = 0;
    NIDT_CUDA_CHECK(cudaGetDevice(&dev));
    NIDT_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const void* kern = pick_topk_kernel(world);
    NIDT_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kSymThreads, 0));
    NIDT_REQUIRE(per_sm >= 1, "sym_saliency_topk: kernel does not fit on an SM");
    // a shard is n/world elements and L2-resident: the passes are bound by the grid-wide and cross-rank synchronisations, whose cost
    // grows with the number of CTAs -- 4 to 48 CTAs
    int64_t want = (n + 16383) / 16384;        // the mask / score copy-outs run over the whole slab on every rank
    if (want < 4) want = 4;
    // NIDT_SAL_REPLICATE=0|1 (default: on for slabs up to 64 MB): select on every rank over the all-gathered scores instead of
    // exchanging three shard histograms -- 2 cross-rank barriers instead of 6; the passes then cover the whole slab, on a wider grid
    int replicate = world >= 2 && n <= ((int64_t)16 << 20);
    if (const char* e = getenv("NIDT_SAL_REPLICATE")) replicate = atoi(e) != 0;
    const int cap = replicate ? sms : 48;
    int grid = (int)(want < cap ? want : cap);
    if (grid > sms * per_sm) grid = sms * per_sm;
    int mode = use_multimem ? kModeMultimem : kModeP2P;
    void* args[] = {(void*)&sp, (void*)&aux, (void*)&n, (void*)&flags, (void*)&mask, (void*)&k, (void*)&res, (void*)&scores_out, (void*)&epoch, (void*)&mode,
                    (void*)&replicate};
    NIDT_CUDA_CHECK(cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(kSymThreads), args, 0, (cudaStream_t)stream));
    return 0;
}

NIDT_API int nidt_sym_saliency_read_result(const void* scratch, float* threshold, float* norm, unsigned int* kept, void* stream) {
    SalResult h;
    NIDT_CUDA_CHECK(cudaMemcpyAsync(&h, scratch, sizeof(h), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    NIDT_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    if (threshold) *threshold = h.threshold;
    if (norm) *norm = h.norm;
    if (kept) *kept = h.kept;
    return 0;
}
