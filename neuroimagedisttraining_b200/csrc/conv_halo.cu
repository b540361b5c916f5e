// 3x3x3 / stride-1 convolution with the input halo RESIDENT in shared memory (tcgen05 + TMA).
//
// conv_tc.cu fetches one box of input positions per filter tap, i.e. every input element travels L2 -> SM up to 27 times;
// ncu shows those kernels pinned at the L2 -> SM bandwidth (~30 B/clk/SM of the ~43 B/clk/SM chip-wide cap) with the
// tensor pipe 20-30 % busy.  Here a CTA keeps three input depth-planes of its row block in shared memory and slides along
// the depth axis: every step loads ONE new plane (TMA, SWIZZLE_128B, zero fill = padding) and issues all 27 taps against
// the resident planes -- a tap is nothing but a different start address of the same UMMA descriptor, because the GEMM M
// rows are the flattened (line, w) positions of the plane with the halo pitch W+2 (rows that fall on the two halo columns
// of a line compute garbage and are dropped by the epilogue).  Measured on B200: the 128-byte swizzle is a function of the
// shared-memory ADDRESS bits, so a descriptor that starts in the middle of a 1024-byte swizzle atom needs no base offset
// (setting the descriptor's base-offset field to (start >> 7) & 7 gives wrong results).  The weight tile of a tap is fetched once per step and reused
// by all T accumulators (T x 128 rows) of the step, so per 128 output rows the L2 -> SM traffic drops from
// 27 x (16 KB + B) to (plane + 27 x B) / T.
//
//   warp 0  : TMA producer (planes into a 3-slot ring keyed by input depth mod 3, weight tiles into a small ring)
//   warp 1  : MMA issuer (one thread), accumulators double-buffered in TMEM (2 x T x BLOCK_N columns)
//   warps 2-9: epilogue: TMEM -> registers -> (+bias) -> bf16 -> row stores; BatchNorm (sum, sum^2) partials accumulate per
//              thread (= per output row) in registers and are reduced across lanes once, when the CTA finishes.
// Work = linear index over (n_tile, sample, row block, depth); each CTA owns one contiguous range.
// dgrad = the same kernel on dY with the flipped / transposed weights and pad' = 2 - pad.
// Reference: nn.Conv3d forward/backward (cuDNN fp32), fedml_api/model/cv/salient_models.py:20-40,152-166.
#include "tc_common.cuh"
#include <stdlib.h>

namespace nidt {
namespace halo {

constexpr int kThreads = 320;      // warp 0 producer, warp 1 MMA, warps 2-9 epilogue
constexpr int kSmemLimit = 227 * 1024;

struct Geom {
    int N, Do, Ho, Wo, Cout;
    int pad, cin_blocks;
    int Wbox;          // Wo + 2: row pitch of a plane in shared memory
    int Lb, HB;        // output lines per row block, row blocks per plane
    int T;             // 128-row accumulators per step = ceil(Lb*Wbox / 128)
    int plane_rows;    // shared-memory rows per (slot, cin block), multiple of 8
    int n_tiles;       // Cout / BLOCK_N
    int total_steps;   // n_tiles * N * HB * Do
    int b_stages;
    int debug;         // bit 1: the MMA warp of CTA 0 prints where its cycles went
};

// halving butterfly: on return lanes hold the sum over the 32 lanes of v[idx], idx = bits 4..1 of the lane (bit 4 -> 8, ..., bit 1 -> 1)
__device__ __forceinline__ float reduce16(float* v, int lane) {
#pragma unroll
    for (int half = 8, bit = 16; half >= 1; half >>= 1, bit >>= 1) {
        const bool up = (lane & bit) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float send = up ? v[i] : v[i + half];
            const float keep = up ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, bit);
        }
    }
    return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

template <int BLOCK_N>
__global__ void __launch_bounds__(kThreads, 1) conv3d_halo_kernel(
    const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_w, const Geom g, __nv_bfloat16* __restrict__ y,
    const float* __restrict__ bias, float* __restrict__ stats_partials) {
    constexpr int kBTileBytes = BLOCK_N * 128;
    constexpr int kChunks = BLOCK_N / 16;
    extern __shared__ __align__(1024) uint8_t smem[];
    if ((tc::smem_u32(smem) & 1023u) != 0u) __trap();
    const int plane_bytes = g.plane_rows * 128;                 // one cin block of one slot
    const int slot_bytes = plane_bytes * g.cin_blocks;
    uint8_t* planes = smem;
    uint8_t* bring = smem + 3 * slot_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(bring + g.b_stages * kBTileBytes);
    uint64_t* plane_full = bars;            // [3]
    uint64_t* plane_empty = bars + 3;       // [3]
    uint64_t* b_full = bars + 6;            // [b_stages]
    uint64_t* b_empty = b_full + 16;        // [b_stages] (<= 16)
    uint64_t* tmem_full = b_empty + 16;     // [2]
    uint64_t* tmem_empty = tmem_full + 2;   // [2]
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    float* bias_s = reinterpret_cast<float*>(tmem_ptr + 4);    // [BLOCK_N], 16-byte aligned

    // warp index through a shuffle: the compiler then knows it is warp-uniform, keeps the producer / MMA schedules in uniform
    // registers and issues UTMALDG / UTCHMMA without a per-instruction lane-election loop (the whole warp walks the schedule,
    // one elected lane issues)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&tm_x);
        tc::tma_prefetch_desc(&tm_w);
        for (int i = 0; i < 3; ++i) { tc::mbar_init(&plane_full[i], 1); tc::mbar_init(&plane_empty[i], 1); }
        for (int i = 0; i < g.b_stages; ++i) { tc::mbar_init(&b_full[i], 1); tc::mbar_init(&b_empty[i], 1); }
        for (int i = 0; i < 2; ++i) { tc::mbar_init(&tmem_full[i], 1); tc::mbar_init(&tmem_empty[i], 256); }
        tc::mbar_fence_init();
    }
    if (warp == 1) tc::tmem_alloc<512>(tmem_ptr);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    // contiguous range of steps owned by this CTA
    const int s_begin = (int)(((int64_t)g.total_steps * blockIdx.x) / gridDim.x);
    const int s_end = (int)(((int64_t)g.total_steps * (blockIdx.x + 1)) / gridDim.x);
    auto decode = [&](int s, int& nt, int& n, int& hb, int& d) {
        d = s % g.Do; s /= g.Do;
        hb = s % g.HB; s /= g.HB;
        n = s % g.N;
        nt = s / g.N;
    };
    const int box_bytes = g.Wbox * (g.Lb + 2) * 128;   // bytes one plane TMA delivers per cin block

    if (warp == 0) {
        const bool leader = tc::elect_one();
        {
            // ===== TMA producer =====
            uint32_t fill_par = 0;      // bit i = parity of the number of loads into slot i
            uint32_t st = 0, ph = 0;
            for (int s = s_begin; s < s_end; ++s) {
                int nt, n, hb, d;
                decode(s, nt, n, hb, d);
                const bool new_col = (s == s_begin) || d == 0;
                const int h_in = hb * g.Lb - g.pad;
                for (int kd = new_col ? 0 : 2; kd < 3; ++kd) {
                    const int slot = (d + kd) % 3;
                    tc::mbar_wait(&plane_empty[slot], ((fill_par >> slot) & 1) ^ 1);
                    fill_par ^= 1u << slot;
                    if (leader) {
                        tc::mbar_expect_tx(&plane_full[slot], (uint32_t)(box_bytes * g.cin_blocks));
                        for (int cb = 0; cb < g.cin_blocks; ++cb)
                            tc::tma_load_5d(&tm_x, &plane_full[slot], planes + slot * slot_bytes + cb * plane_bytes, cb * 64, -g.pad, h_in, d - g.pad + kd, n);
                    }
                }
                for (int kb = 0; kb < 27 * g.cin_blocks; ++kb) {       // weight tiles in (tap, cin block) order
                    tc::mbar_wait(&b_empty[st], ph ^ 1);
                    if (leader) {
                        tc::mbar_expect_tx(&b_full[st], kBTileBytes);
                        tc::tma_load_2d(&tm_w, &b_full[st], bring + st * kBTileBytes, kb * 64, nt * BLOCK_N);
                    }
                    if (++st == (uint32_t)g.b_stages) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        const bool leader = tc::elect_one();
        {
            // ===== MMA issuer =====
            constexpr uint32_t idesc = tc::make_idesc_bf16(128, BLOCK_N, 0, 0);
            constexpr uint32_t kDescHi = tc::desc_hi_sw128(1024);
            uint32_t fill_par = 0;
            uint32_t st = 0, ph = 0, local = 0;      // weight-ring stage / phase
            const uint32_t planes_lo = tc::desc_lo_sw128(tc::smem_u32(planes)), bring_lo = tc::desc_lo_sw128(tc::smem_u32(bring));
            long long w_b = 0, w_plane = 0, w_tmem = 0, t_all = clock64(), t0;
            const bool prof = (g.debug & 2) != 0;
            const int64_t iters_total = (int64_t)(s_end - s_begin) * 27 * g.cin_blocks;      // weight tiles this CTA consumes
            int64_t iter = 0;
            bool waited = false;
            for (int s = s_begin; s < s_end; ++s, ++local) {
                int nt, n, hb, d;
                decode(s, nt, n, hb, d);
                const bool new_col = (s == s_begin) || d == 0;
                const bool last_of_col = (s + 1 == s_end) || (d + 1 == g.Do);
                const int lines = min(g.Lb, g.Ho - hb * g.Lb);
                const int T = (lines * g.Wbox + 127) >> 7;
                const int buf = local & 1;
                if (prof) t0 = clock64();
                tc::mbar_wait(&tmem_empty[buf], ((local >> 1) & 1) ^ 1);
                if (prof) w_tmem += clock64() - t0;
                tc::tc_fence_after();
                const uint32_t d_tmem0 = tmem_base + buf * g.T * BLOCK_N;
                uint32_t acc = 0;
                for (int kd = 0; kd < 3; ++kd) {
                    const int slot = (d + kd) % 3;
                    if (new_col || kd == 2) {
                        if (prof) t0 = clock64();
                        tc::mbar_wait(&plane_full[slot], (fill_par >> slot) & 1);
                        if (prof) w_plane += clock64() - t0;
                        fill_par ^= 1u << slot;
                        tc::tc_fence_after();
                    }
                    // no divisions / 64-bit work in this loop: with one warp per scheduler every scalar instruction costs
                    // ~5-10 cycles of issue latency, and a tap is only 4*T MMAs (~50 cycles each) long
                    const uint32_t slot_lo = planes_lo + slot * (slot_bytes >> 4);
                    uint32_t tap_lo = slot_lo;                   // + (kh*Wbox + kw) rows of 128 B = 8 descriptor units each
                    for (int kh = 0; kh < 3; ++kh, tap_lo += (g.Wbox - 3) * 8)
                        for (int kw = 0; kw < 3; ++kw, tap_lo += 8) {
                            uint32_t a_lo = tap_lo;
                            for (int cb = 0; cb < g.cin_blocks; ++cb, a_lo += plane_bytes >> 4, ++iter) {
                                if (!waited) {      // first tile of the CTA / of a new column: nothing could be checked ahead
                                    if (prof) t0 = clock64();
                                    tc::mbar_wait(&b_full[st], ph);
                                    if (prof) w_b += clock64() - t0;
                                    tc::tc_fence_after();
                                }
                                const uint32_t b_lo = bring_lo + st * (kBTileBytes >> 4);
                                uint32_t st1 = st + 1, ph1 = ph;
                                if (st1 == (uint32_t)g.b_stages) { st1 = 0; ph1 ^= 1; }
                                if (leader) {
                                    tc::umma_bf16_lohi(d_tmem0, a_lo, b_lo, kDescHi, idesc, acc);
                                    tc::umma_bf16_lohi(d_tmem0, a_lo + 2, b_lo + 2, kDescHi, idesc, 1);
                                }
                                // the NEXT weight tile's barrier is checked in the middle of this tap's MMAs (in front of a group it
                                // would idle the tensor pipe for ~100 cycles: scripts/bench_native/mma_rate.cu, modes 2 / 3)
                                // (not across a column change: the producer loads the new column's planes first, and the last of
                                // their slots is only released by the commit that follows this very tap)
                                waited = (iter + 1 < iters_total) && !(last_of_col && kd == 2 && kh == 2 && kw == 2 && cb + 1 == g.cin_blocks);
                                if (waited) {
                                    if (prof) t0 = clock64();
                                    tc::mbar_wait(&b_full[st1], ph1);
                                    if (prof) w_b += clock64() - t0;
                                    tc::tc_fence_after();
                                }
                                if (leader) {
                                    tc::umma_bf16_lohi(d_tmem0, a_lo + 4, b_lo + 4, kDescHi, idesc, 1);
                                    tc::umma_bf16_lohi(d_tmem0, a_lo + 6, b_lo + 6, kDescHi, idesc, 1);
                                    for (int t = 1; t < T; ++t) {        // tile t starts 128 rows = 16 KB further: +1024 in the low word
                                        const uint32_t d_tmem = d_tmem0 + t * BLOCK_N;
                                        tc::umma_bf16_lohi(d_tmem, a_lo + t * 1024, b_lo, kDescHi, idesc, acc);
                                        tc::umma_bf16_lohi(d_tmem, a_lo + t * 1024 + 2, b_lo + 2, kDescHi, idesc, 1);
                                        tc::umma_bf16_lohi(d_tmem, a_lo + t * 1024 + 4, b_lo + 4, kDescHi, idesc, 1);
                                        tc::umma_bf16_lohi(d_tmem, a_lo + t * 1024 + 6, b_lo + 6, kDescHi, idesc, 1);
                                    }
                                    tc::umma_commit(&b_empty[st]);
                                }
                                acc = 1;
                                st = st1; ph = ph1;
                            }
                        }
                    if (leader && (kd == 0 || last_of_col)) tc::umma_commit(&plane_empty[slot]);
                }
                if (leader) tc::umma_commit(&tmem_full[buf]);
            }
            if (prof && leader && (blockIdx.x == 0 || blockIdx.x == 100))
                printf("halo cta %d: %d steps, %lld cycles; waiting for weights %lld, planes %lld, epilogue %lld\n", blockIdx.x, s_end - s_begin,
                       clock64() - t_all, w_b, w_plane, w_tmem);
        }
    } else {
        // ===== epilogue: warps 2..9.  TMEM lane quarter = warp % 4; the two warps of a quarter split the columns =====
        // Each thread owns one output row per tile and keeps per-channel (sum, sum^2) of ITS rows in registers; the cross-lane
        // reduction happens once per CTA (a per-tile butterfly made these warps the critical path: ncu, 30 k instructions/warp).
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        constexpr int kCols = BLOCK_N / 2;           // columns per warp
        constexpr int kWChunks = kCols / 16;
        float s1[kCols], s2[kCols];
#pragma unroll
        for (int i = 0; i < kCols; ++i) s1[i] = s2[i] = 0.f;
        const int ch_in_chunk = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
        auto flush = [&](int nt) {
            if (stats_partials != nullptr) {
                const size_t prow = (size_t)blockIdx.x * 4 + q;
#pragma unroll
                for (int c = 0; c < kWChunks; ++c) {
                    const float r1 = reduce16(&s1[c * 16], lane);
                    const float r2 = reduce16(&s2[c * 16], lane);
                    if ((lane & 1) == 0) {
                        const int ch = nt * BLOCK_N + half * kCols + c * 16 + ch_in_chunk;
                        stats_partials[(prow * 2 + 0) * g.Cout + ch] = r1;
                        stats_partials[(prow * 2 + 1) * g.Cout + ch] = r2;
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < kCols; ++i) s1[i] = s2[i] = 0.f;
        };
        uint32_t local = 0;
        int cur_nt = -1;
        for (int s = s_begin; s < s_end; ++s, ++local) {
            int nt, n, hb, d;
            decode(s, nt, n, hb, d);
            if (nt != cur_nt) {
                if (cur_nt >= 0) flush(cur_nt);
                cur_nt = nt;
                if (bias != nullptr) {      // this CTA's bias slice -> shared memory (read back as broadcast float4)
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    if (threadIdx.x - 64 < BLOCK_N) bias_s[threadIdx.x - 64] = bias[nt * BLOCK_N + threadIdx.x - 64];
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                }
            }
            const int h0 = hb * g.Lb;
            const int lines = min(g.Lb, g.Ho - h0);
            const int T = (lines * g.Wbox + 127) >> 7;
            const int buf = local & 1;
            tc::mbar_wait(&tmem_full[buf], (local >> 1) & 1);
            tc::tc_fence_after();
            for (int t = 0; t < T; ++t) {
                const int f = t * 128 + row;
                const int line = f / g.Wbox, w = f - line * g.Wbox;
                const bool valid = (w < g.Wo) && (line < lines);
                __nv_bfloat16* out = y + ((((size_t)n * g.Do + d) * g.Ho + h0 + line) * g.Wo + w) * g.Cout + nt * BLOCK_N + half * kCols;
                const uint32_t taddr = tmem_base + (buf * g.T + t) * BLOCK_N + half * kCols + ((uint32_t)(q * 32) << 16);
#pragma unroll
                for (int c = 0; c < kWChunks; ++c) {
                    float v[16];
                    tc::tmem_ld16(taddr + c * 16, v);
                    if (bias != nullptr) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 bb = *reinterpret_cast<const float4*>(bias_s + half * kCols + c * 16 + 4 * i);
                            v[4 * i] += bb.x; v[4 * i + 1] += bb.y; v[4 * i + 2] += bb.z; v[4 * i + 3] += bb.w;
                        }
                    }
                    uint4 o[2];
                    __nv_bfloat162* ob = reinterpret_cast<__nv_bfloat162*>(o);
#pragma unroll
                    for (int i = 0; i < 8; ++i) ob[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
                    if (valid) {
                        *reinterpret_cast<uint4*>(out + c * 16) = o[0];
                        *reinterpret_cast<uint4*>(out + c * 16 + 8) = o[1];
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float2 r = __bfloat1622float2(ob[i]);     // statistics of the values actually stored
                        const float x0 = valid ? r.x : 0.f, x1 = valid ? r.y : 0.f;   // (garbage rows may hold NaN: select, not multiply)
                        s1[c * 16 + 2 * i] += x0;
                        s1[c * 16 + 2 * i + 1] += x1;
                        s2[c * 16 + 2 * i] = fmaf(x0, x0, s2[c * 16 + 2 * i]);
                        s2[c * 16 + 2 * i + 1] = fmaf(x1, x1, s2[c * 16 + 2 * i + 1]);
                    }
                }
            }
            tc::tc_fence_before();
            tc::mbar_arrive(&tmem_empty[buf]);
        }
        if (cur_nt >= 0) flush(cur_nt);
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc<512>(tmem_base);
    }
}

// picks T / Lb / stages for a shape; returns the dynamic shared-memory size or 0 when the shape does not fit
static size_t plan(Geom& g, int block_n) {
    int t_max = 256 / block_n;     // 2 x T x BLOCK_N <= 512 TMEM columns
    int max_stages = 6;
    if (const char* e = getenv("NIDT_HALO_T")) t_max = atoi(e) < t_max ? atoi(e) : t_max;          // tuning knobs
    if (const char* e = getenv("NIDT_HALO_STAGES")) max_stages = atoi(e) > 16 ? 16 : atoi(e);
    for (int T = t_max; T >= 1; --T) {
        int Lb = (T * 128) / g.Wbox;
        if (Lb < 1) continue;
        if (Lb > g.Ho) Lb = g.Ho;
        const int HB = (g.Ho + Lb - 1) / Lb;
        Lb = (g.Ho + HB - 1) / HB;                       // even out the row blocks
        const int Teff = (Lb * g.Wbox + 127) / 128;
        const int plane_rows = (Teff * 128 + 2 * g.Wbox + 2 + 7) & ~7;
        for (int stages = max_stages; stages >= 3; --stages) {
            const size_t need = (size_t)3 * g.cin_blocks * plane_rows * 128 + (size_t)stages * block_n * 128 + 1024;
            if (need + 1024 <= (size_t)kSmemLimit) {
                g.Lb = Lb; g.HB = HB; g.T = Teff; g.plane_rows = plane_rows; g.b_stages = stages;
                return need;
            }
        }
    }
    return 0;
}

static int fill_geom(Geom& g, int N, int Do, int Ho, int Wo, int Cin, int Cout, int pad, int block_n) {
    g.N = N; g.Do = Do; g.Ho = Ho; g.Wo = Wo; g.Cout = Cout; g.pad = pad; g.cin_blocks = Cin / 64;
    g.Wbox = Wo + 2;
    g.n_tiles = Cout / block_n;
    return 0;
}

}  // namespace halo
}  // namespace nidt
using namespace nidt;

NIDT_API int nidt_conv3d_halo_stats_rows() { return 4 * kNumSMs; }

// 1 when nidt_conv3d_halo_fprop can run the shape (k = 3, stride 1 are implied)
NIDT_API int nidt_conv3d_halo_supported(int N, int Do, int Ho, int Wo, int Cin, int Cout, int pad, int block_n) {
    if (Cin % 64 || (block_n != 64 && block_n != 128) || Cout % block_n || pad < 0 || pad > 2 || Wo + 2 > 128 || Wo < 1) return 0;
    halo::Geom g;
    halo::fill_geom(g, N, Do, Ho, Wo, Cin, Cout, pad, block_n);
    return halo::plan(g, block_n) != 0;
}

// x: bf16 [N][Di][Hi][Wi][Cin]; w: bf16 [Cout][27][Cin] (nidt_pack_conv_weights); y: bf16 [N][Do][Ho][Wo][Cout];
// stats_partials: null or fp32 [nidt_conv3d_halo_stats_rows()][2][Cout] (zeroed here, then filled).
// debug_flags: bit 1 = the MMA warps of two CTAs print a cycle breakdown (tuning aid); otherwise 0.
NIDT_API int nidt_conv3d_halo_fprop(const void* x, const void* w, const float* bias, void* y, float* stats_partials, int N, int Di, int Hi, int Wi,
                                    int Cin, int Do, int Ho, int Wo, int Cout, int pad, int block_n, int debug_flags, void* stream) {
    NIDT_REQUIRE(nidt_conv3d_halo_supported(N, Do, Ho, Wo, Cin, Cout, pad, block_n), "conv3d_halo: unsupported shape (Cin=%d Cout=%d Wo=%d pad=%d block_n=%d)",
                 Cin, Cout, Wo, pad, block_n);
    NIDT_REQUIRE(Do == Di + 2 * pad - 2 && Ho == Hi + 2 * pad - 2 && Wo == Wi + 2 * pad - 2, "conv3d_halo: output extents do not match a 3x3x3 stride-1 conv");
    halo::Geom g;
    halo::fill_geom(g, N, Do, Ho, Wo, Cin, Cout, pad, block_n);
    const size_t smem = halo::plan(g, block_n) + 1024;
    g.debug = debug_flags;
    g.total_steps = g.n_tiles * N * g.HB * Do;
    CUtensorMap tx, tw;
    if (make_tmap_act5d(&tx, x, N, Di, Hi, Wi, Cin, g.Wbox, g.Lb + 2, 1, 1, 1)) return 1;
    if (make_tmap_mat2d(&tw, w, Cout, (int64_t)27 * Cin, block_n)) return 1;
    cudaStream_t s = (cudaStream_t)stream;
    const int grid = g.total_steps < kNumSMs ? g.total_steps : kNumSMs;
    if (stats_partials != nullptr)
        NIDT_CUDA_CHECK(cudaMemsetAsync(stats_partials, 0, (size_t)nidt_conv3d_halo_stats_rows() * 2 * Cout * sizeof(float), s));
    if (block_n == 128) {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(halo::conv3d_halo_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        halo::conv3d_halo_kernel<128><<<grid, halo::kThreads, smem, s>>>(tx, tw, g, (__nv_bfloat16*)y, bias, stats_partials);
    } else {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(halo::conv3d_halo_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        halo::conv3d_halo_kernel<64><<<grid, halo::kThreads, smem, s>>>(tx, tw, g, (__nv_bfloat16*)y, bias, stats_partials);
    }
    NIDT_LAUNCH_CHECK();
    return 0;
}
