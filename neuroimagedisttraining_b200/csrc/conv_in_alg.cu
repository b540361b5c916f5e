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
// (gram_plane_kernel + gram_reduce_kernel, 0.33 ms measured) and kept on the device keyed by the sample's id (ops/engine_b200.py: GramCache);
// a step adds up the matrices of its 16 samples (gram_sum_kernel).  What is left per step is S (alg_sparse_kernel: 8.5 M patch
// reads instead of 4 M positions x 64 channels x 216 taps on the tensor cores and two 506 MB passes) and a 64 x 125 x 125 fp64
// product (alg_combine_kernel).  The bias gradient is exactly zero (BatchNorm subtracts the batch mean right after the conv).
//
// x~ = bf16(u8 * scale) and W~ = bf16(W) are the operands the forward kernel (conv_in_s2d.cu) multiplies, so this is the gradient of
// the function the forward pass computes; y enters un-rounded, which the dY-based kernels cannot do.
// Reference op: autograd through Conv3d + BatchNorm3d + ReLU + MaxPool3d, fedml_api/model/cv/salient_models.py:147-150 (cuDNN).
#include "common.cuh"
#include "../../include/nidt_b200.h"
#include <stdlib.h>
#include <type_traits>

namespace nidt {
namespace alg {

constexpr int kTaps = 125;        // 5 x 5 x 5
constexpr int kTP = 128;          // padded: tap 125 = constant 1 (gives sx and the position count), 126 / 127 = 0
constexpr int kCout = 64;
constexpr int kMaxList = 64;

struct IntList { int n; int v[kMaxList]; };

// acc += x * g with x, g bf16 halves of 32-bit registers: sm_100's mixed-precision FMA (fma.rn.f32.bf16 -> FHFMA.BF16 with .H0 / .H1
// operand selectors) -- the product of two bf16 is exact in fp32, so this equals fmaf(float(x), float(g), acc) without the unpacking
// instructions.  `gw`: the gradient in the HIGH half.
__device__ __forceinline__ void fma_pair(float& acc_lo, float& acc_hi, uint32_t pair, uint32_t gw) {
    asm("{\n .reg .b16 xl, xh, gl, gh;\n mov.b32 {xl, xh}, %2;\n mov.b32 {gl, gh}, %3;\n fma.rn.f32.bf16 %0, xl, gh, %0;\n fma.rn.f32.bf16 %1, xh, gh, %1;\n}\n"
        : "+f"(acc_lo), "+f"(acc_hi) : "r"(pair), "r"(gw));
}
__device__ __forceinline__ void fma_lo(float& acc_lo, uint32_t pair, uint32_t gw) {
    asm("{\n .reg .b16 xl, xh, gl, gh;\n mov.b32 {xl, xh}, %1;\n mov.b32 {gl, gh}, %2;\n fma.rn.f32.bf16 %0, xl, gh, %0;\n}\n" : "+f"(acc_lo) : "r"(pair), "r"(gw));
}

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
// Unit of work: one row of up to kG pooling windows along W, (n, dp, hp, chunk).  The 9 planes x 9 rows x (6g+3) input columns
// the row covers are contiguous byte runs in global memory (coalesced loads) and go to shared memory as x~ in bf16 PAIRS (x~ is
// bf16-exact; one 32-bit load feeds two taps), double-buffered, together with every (window, channel)'s pooled gradient and the
// word offset of its winning patch.  Thread (c = tid % 64, tg = tid / 64) then adds  da[w][c] * x~[winner patch of (w,c)][t]  for
// its 6-7 of the 25 (td,th) tap rows (5 taps each) into registers.
// The loop is bound by shared-memory loads (75 per patch), so the layout is made conflict-free: a winner (ad,ah,aw) moves the patch
// by 2*ad planes, 2*ah rows, 2*aw columns; planes and rows are stored EVEN ONES FIRST (slot = i/2 + 5*(i%2)), which turns those
// into +ad plane slots, +ah row slots, +aw words, and with row pitch = 3, plane pitch = 9 (mod 32 words) the 27 possible winners
// of a warp's 32 channels read 27 different banks.
constexpr int kG = 19;                      // windows per unit
constexpr int kRowW = 67;                   // words per strip row  (>= ceil((6*kG + 3) / 2) = 59;  = 3 mod 32)
constexpr int kPlaneW = 617;                // words per plane      (>= 9 * kRowW;                  = 9 mod 32)
constexpr int kStripW = 9 * kPlaneW;        // 5553 words
__host__ __device__ constexpr int slot9(int i) { return (i >> 1) + 5 * (i & 1); }
struct __align__(16) StripBuf {
    uint32_t nb[kStripW + 3];
    uint32_t pw[kG * kCout];                // per (window, channel): low half = word offset of the winning patch inside the strip (without
                                            // 3*wi; 0xFFFF = gated by ReLU), high half = the pooled gradient's bf16 bits
};
constexpr int kSparseSmem = 2 * (int)sizeof(StripBuf);
constexpr int kSparseRows = 7;              // tap rows per thread: 7, 6, 6, 6

template <int kCtasPerSM>
__global__ void __launch_bounds__(256, kCtasPerSM) alg_sparse_kernel(const unsigned char* __restrict__ x, const unsigned char* __restrict__ amax,
                                                           const __nv_bfloat16* __restrict__ da, int N, int Di, int Hi, int Wi, int Dp, int Hp, int Wp,
                                                           float scale, float* __restrict__ partials) {
    extern __shared__ __align__(16) unsigned char sp_smem[];
    StripBuf* bufs = reinterpret_cast<StripBuf*>(sp_smem);
    const int tid = threadIdx.x, c = tid & 63, tg = tid >> 6;
    const int r0 = tg == 0 ? 0 : 1 + 6 * tg, nr = tg == 0 ? 7 : 6;
    int roff[kSparseRows];
#pragma unroll
    for (int i = 0; i < kSparseRows; ++i) {
        const int r = r0 + (i < nr ? i : 0);
        roff[i] = slot9(r / 5) * kPlaneW + slot9(r % 5) * kRowW;
    }
    float acc[kSparseRows][5];
#pragma unroll
    for (int i = 0; i < kSparseRows; ++i)
#pragma unroll
        for (int j = 0; j < 5; ++j) acc[i][j] = 0.f;
    const int chunks = (Wp + kG - 1) / kG;
    const int units = N * Dp * Hp * chunks;
    // The next unit's bytes are loaded into REGISTERS before the current unit is processed and converted / stored after it, so
    // the global-memory latency hides behind the compute loop.  Thread (q = tid / 64, j = tid % 64) fetches word j (two bytes) of
    // rows q and q + 4 of every plane, q = 0 also of row 8: four bytes per register, all offsets compile-time multiples of Wi, Hi*Wi.
    constexpr int kWd = (kG * kCout + 255) / 256;           // 5
    uint32_t va[9];                                         // plane z: bytes of row q (low half) and row q + 4 (high half)
    uint32_t vb[5];                                         // q == 0: bytes of row 8 of planes 2m (low half) and 2m + 1 (high half)
    uint32_t pw[kWd];                                       // winning tap (low half) | pooled gradient bf16 bits (high half)
    const int lq = tid >> 6, lj = tid & 63;
    const int HiWi = Hi * Wi;
    const int dst_q = slot9(lq) * kRowW + lj;               // row q -> its slot; row q + 4 is two slots further, row 8 is slot 4
    // unit walker without per-unit integer divisions: units advance by gridDim.x, so (ch, hp, dp, n) advance by fixed deltas with carries
    int u_ch, u_hp, u_dp, u_n, s_ch, s_hp, s_dp, s_n;
    {
        const int u0 = blockIdx.x, st = gridDim.x;
        u_ch = u0 % chunks; u_hp = (u0 / chunks) % Hp; u_dp = (u0 / (chunks * Hp)) % Dp; u_n = u0 / (chunks * Hp * Dp);
        s_ch = st % chunks; s_hp = (st / chunks) % Hp; s_dp = (st / (chunks * Hp)) % Dp; s_n = st / (chunks * Hp * Dp);
    }
    auto advance = [&]() {
        u_ch += s_ch;
        int cy = 0;
        if (u_ch >= chunks) { u_ch -= chunks; cy = 1; }
        u_hp += s_hp + cy; cy = 0;
        if (u_hp >= Hp) { u_hp -= Hp; cy = 1; }
        u_dp += s_dp + cy; cy = 0;
        if (u_dp >= Dp) { u_dp -= Dp; cy = 1; }
        u_n += s_n + cy;
    };
    auto issue = [&]() {                    // loads of the unit (u_ch, u_hp, u_dp, u_n)
        const int w0 = u_ch * kG, g = min(kG, Wp - w0);
        const int ncols = 6 * g + 3;
        const unsigned char* src = x + (((size_t)u_n * Di + 6 * u_dp) * Hi + 6 * u_hp) * Wi + 6 * w0 + 2 * lj + lq * Wi;
        const bool in0 = 2 * lj < ncols, in1 = 2 * lj + 1 < ncols;
        auto two = [&](const unsigned char* s2) -> uint32_t {
            const uint32_t b0 = in0 ? s2[0] : 0u, b1 = in1 ? s2[1] : 0u;
            return b0 | (b1 << 8);
        };
#pragma unroll
        for (int z = 0; z < 9; ++z) va[z] = two(src + z * HiWi) | (two(src + z * HiWi + 4 * Wi) << 16);
        if (lq == 0) {
#pragma unroll
            for (int m = 0; m < 5; ++m) {
                vb[m] = two(src + (2 * m) * HiWi + 8 * Wi);
                if (2 * m + 1 < 9) vb[m] |= two(src + (2 * m + 1) * HiWi + 8 * Wi) << 16;
            }
        }
        const size_t wbase = ((((size_t)u_n * Dp + u_dp) * Hp + u_hp) * Wp + w0) * kCout;
#pragma unroll
        for (int k = 0; k < kWd; ++k) {
            const int i = tid + 256 * k;
            const bool ok = i < g * kCout;
            const unsigned int am = ok ? amax[wbase + i] : 255u;          // (used in commit: no wait for the load here)
            const __nv_bfloat16 dv = ok ? da[wbase + i] : __float2bfloat16(0.f);
            pw[k] = am | ((uint32_t)__bfloat16_as_ushort(dv) << 16);
        }
    };
    auto commit = [&](StripBuf& sb) {
        auto put = [&](int word, uint32_t two_bytes) {
            __nv_bfloat162 v = __floats2bfloat162_rn((float)(two_bytes & 255u) * scale, (float)((two_bytes >> 8) & 255u) * scale);
            sb.nb[word] = *reinterpret_cast<uint32_t*>(&v);
        };
#pragma unroll
        for (int z = 0; z < 9; ++z) {
            put(slot9(z) * kPlaneW + dst_q, va[z]);
            put(slot9(z) * kPlaneW + dst_q + 2 * kRowW, va[z] >> 16);
        }
        if (lq == 0) {
#pragma unroll
            for (int m = 0; m < 5; ++m) {
                put(slot9(2 * m) * kPlaneW + 4 * kRowW + lj, vb[m]);
                if (2 * m + 1 < 9) put(slot9(2 * m + 1) * kPlaneW + 4 * kRowW + lj, vb[m] >> 16);
            }
        }
#pragma unroll
        for (int k = 0; k < kWd; ++k) {
            const int i = tid + 256 * k;
            const uint32_t am = pw[k] & 0xFFFFu;
            const uint32_t po = am == 255u ? 0xFFFFu : (am / 9) * kPlaneW + ((am / 3) % 3) * kRowW + am % 3;
            if (i < kG * kCout) sb.pw[i] = (pw[k] & 0xFFFF0000u) | po;
        }
    };
    int u = blockIdx.x, it = 0;
    if (u < units) { issue(); commit(bufs[0]); }
    for (; u < units; u += gridDim.x, ++it) {
        __syncthreads();                                   // bufs[it & 1] is complete; everybody is done with bufs[(it + 1) & 1]
        const int g = min(kG, Wp - u_ch * kG);             // windows of the unit being processed
        const int un = u + gridDim.x;
        advance();
        if (un < units) issue();
        const StripBuf& sb = bufs[it & 1];
        for (int wi = 0; wi < g; ++wi) {
            const uint32_t w = sb.pw[wi * kCout + c];
            if ((w & 0xFFFFu) == 0xFFFFu) continue;
            const uint32_t* p = sb.nb + (w & 0xFFFFu) + 3 * wi;
#pragma unroll
            for (int i = 0; i < kSparseRows; ++i) {
                if (i < nr) {
                    const uint32_t a0 = p[roff[i]], a1 = p[roff[i] + 1], a2 = p[roff[i] + 2];
                    fma_pair(acc[i][0], acc[i][1], a0, w);
                    fma_pair(acc[i][2], acc[i][3], a1, w);
                    fma_lo(acc[i][4], a2, w);
                }
            }
        }
        if (un < units) commit(bufs[(it + 1) & 1]);
    }
    float* out = partials + ((size_t)blockIdx.x * kCout + c) * kTP;
#pragma unroll
    for (int i = 0; i < kSparseRows; ++i)
        if (i < nr) {
#pragma unroll
            for (int j = 0; j < 5; ++j) out[(r0 + i) * 5 + j] = acc[i][j];
        }
}

// ------------------------------------------------------------------------------------------------ combine
// grid 64 (channel), 512 threads = 4 groups x 128 taps: every group adds a quarter of the sparse partials and of the W~ G product
// (fp64: the bracket is a difference of sums over 4 M positions), then the groups are added.  dW as in the file header.
// Optional SNIP epilogue: scores += sal_scale * |dW * w|.
__global__ void __launch_bounds__(512) alg_combine_kernel(const float* __restrict__ partials, int n_part, const double* __restrict__ gb,
                                                         const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ bn_scale,
                                                         const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ c12,
                                                         float* __restrict__ grad_w, float* __restrict__ grad_b, const float* __restrict__ sal_w,
                                                         float* __restrict__ sal_scores, float sal_scale) {
    __shared__ double wq[kTP];
    __shared__ double red[2][4][kTP];
    const int c = blockIdx.x, t = threadIdx.x & 127, grp = threadIdx.x >> 7;
    if (grp == 0) wq[t] = t < kTaps ? (double)__bfloat162float(__float2bfloat16(W[c * kTaps + t])) : 0.0;
    __syncthreads();
    double s0 = 0.0, s1 = 0.0;
    int b = grp;
    for (; b + 4 < n_part; b += 8) {
        s0 += (double)partials[((size_t)b * kCout + c) * kTP + t];
        s1 += (double)partials[((size_t)(b + 4) * kCout + c) * kTP + t];
    }
    if (b < n_part) s0 += (double)partials[((size_t)b * kCout + c) * kTP + t];
    double w0 = 0.0, w1 = 0.0;
    int u = grp;
    for (; u + 4 < kTaps; u += 8) {
        w0 = fma(wq[u], gb[u * kTP + t], w0);
        w1 = fma(wq[u + 4], gb[(u + 4) * kTP + t], w1);
    }
    if (u < kTaps) w0 = fma(wq[u], gb[u * kTP + t], w0);
    red[0][grp][t] = s0 + s1;
    red[1][grp][t] = w0 + w1;
    __syncthreads();
    if (grp != 0) return;
    const double S = red[0][0][t] + red[0][1][t] + red[0][2][t] + red[0][3][t];
    const double wg = red[1][0][t] + red[1][1][t] + red[1][2][t] + red[1][3][t];
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
static int alg_ctas_per_sm() {
    static const int v = [] { const char* e = getenv("NIDT_ALG_CTAS"); const int n = e ? atoi(e) : 2; return n == 3 ? 3 : 2; }();
    return v;
}
NIDT_API int nidt_conv_in_alg_grid(void) { return 3 * kNumSMs; }      // upper bound (scratch sizing); the launch uses alg_ctas_per_sm() * SMs
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
    const int units = N * Dp * Hp * ((Wp + alg::kG - 1) / alg::kG);
    const int per_sm = alg_ctas_per_sm();
    int grid = per_sm * kNumSMs;
    if (grid > units) grid = units;
    cudaStream_t st = (cudaStream_t)stream;
    static const int dbg = [] { const char* e = getenv("NIDT_ALG_DEBUG"); return e ? atoi(e) : 0; }();      // timing aid: 1 = no sparse pass, 2 = no combine
    static bool attr_set = false;
    if (!attr_set) {
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(alg::alg_sparse_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, alg::kSparseSmem));
        NIDT_CUDA_CHECK(cudaFuncSetAttribute(alg::alg_sparse_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, alg::kSparseSmem));
        attr_set = true;
    }
    if (!(dbg & 1)) {
        if (per_sm == 2)
            alg::alg_sparse_kernel<2><<<grid, 256, alg::kSparseSmem, st>>>((const unsigned char*)x, (const unsigned char*)amax, (const __nv_bfloat16*)da,
                                                                           N, Di, Hi, Wi, Dp, Hp, Wp, scale, scratch);
        else
            alg::alg_sparse_kernel<3><<<grid, 256, alg::kSparseSmem, st>>>((const unsigned char*)x, (const unsigned char*)amax, (const __nv_bfloat16*)da,
                                                                           N, Di, Hi, Wi, Dp, Hp, Wp, scale, scratch);
        NIDT_LAUNCH_CHECK();
    }
    if (!(dbg & 2))
        alg::alg_combine_kernel<<<alg::kCout, 512, 0, st>>>(scratch, grid, gb, W, bias, bn_scale, mean, invstd, c12, grad_w, grad_b, sal_w, sal_scores,
                                                           sal_scale);
    NIDT_LAUNCH_CHECK();
    return 0;
}
