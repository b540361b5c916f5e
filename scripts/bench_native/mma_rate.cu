// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, bf16, M=128, K=16, SWIZZLE_128B K-major operands resident in
// shared memory) from one elected thread, for N = 64 / 128 / 256.  Prints cycles per MMA.  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I neuroimagedisttraining_b200/csrc scripts/bench_native/mma_rate.cu \
//        neuroimagedisttraining_b200/build/obj/common_impl.o -o /tmp/mma_rate   (see scripts/bench_native/run.sh)
#include "tc_common.cuh"
#include <cstdio>
#include <cstdarg>
void nidt_set_error(const char* fmt, ...) { va_list a; va_start(a, fmt); vfprintf(stderr, fmt, a); va_end(a); }
using namespace nidt;

template <int N, int kDistinctA>
__global__ void __launch_bounds__(128, 1) rate_kernel(long long* out, int iters, int a_shift_rows) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_ptr;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    for (int i = threadIdx.x; i < (80 + 32) * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { tc::mbar_init(&bar, 1); tc::mbar_fence_init(); }
    if (warp == 0) tc::tmem_alloc<512>(&tmem_ptr);
    tc::fence_proxy_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = tmem_ptr;
    if (warp == 1) {
        const bool leader = tc::elect_one();
        constexpr uint32_t idesc = tc::make_idesc_bf16(128, N, 0, 0);
        const uint32_t a_addr = tc::smem_u32(smem) + a_shift_rows * 128, b_addr = tc::smem_u32(smem) + 80 * 1024;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint64_t da = tc::make_smem_desc_sw128(a_addr + (kDistinctA ? ((it & 3) * 16384) : 0) + k * 32, 16, 1024);
                const uint64_t db = tc::make_smem_desc_sw128(b_addr + k * 32, 16, 1024);
                if (leader) tc::umma_bf16(tmem, da, db, idesc, 1);
            }
        }
        if (leader) tc::umma_commit(&bar);
        tc::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (leader) out[blockIdx.x] = t1 - t0;
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc::tc_fence_after(); tc::tmem_dealloc<512>(tmem); }
}


// groups of kGroup MMAs; mode 1: tcgen05.commit to a barrier after each group (nobody waits on it);
// mode 2: additionally a try_wait on an already-completed barrier + tcgen05.fence::after_thread_sync before each group
template <int N, int kGroup, int kMode>
__global__ void __launch_bounds__(128, 1) group_kernel(long long* out, int groups) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar, bars[8], done_bar;
    __shared__ uint32_t tmem_ptr;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    for (int i = threadIdx.x; i < (80 + 32) * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) {
        tc::mbar_init(&bar, 1); tc::mbar_init(&done_bar, 1);
        for (int i = 0; i < 8; ++i) tc::mbar_init(&bars[i], 1);
        tc::mbar_fence_init();
        tc::mbar_arrive(&done_bar);       // phase 0 complete: waits with parity 0 pass immediately
    }
    if (warp == 0) tc::tmem_alloc<512>(&tmem_ptr);
    tc::fence_proxy_async_smem();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = tmem_ptr;
    if (warp == 1) {
        const bool leader = tc::elect_one();
        constexpr uint32_t idesc = tc::make_idesc_bf16(128, N, 0, 0);
        constexpr uint32_t hi = tc::desc_hi_sw128(1024);
        const uint32_t a_lo = tc::desc_lo_sw128(tc::smem_u32(smem)), b_lo = tc::desc_lo_sw128(tc::smem_u32(smem) + 80 * 1024);
        long long t0 = clock64();
        for (int gi = 0; gi < groups; ++gi) {
            if (kMode == 2) { tc::mbar_wait(&done_bar, 0); tc::tc_fence_after(); }
            if (kMode == 3) {       // the wait for the NEXT group's operands sits in the middle of this group's MMAs
                if (leader) {
#pragma unroll
                    for (int m = 0; m < kGroup / 2; ++m) tc::umma_bf16_lohi(tmem + (m >> 2) * N % 256, a_lo + (m >> 2) * 1024 + (m & 3) * 2, b_lo + (m & 3) * 2, hi, idesc, 1);
                }
                tc::mbar_wait(&done_bar, 0); tc::tc_fence_after();
                if (leader) {
#pragma unroll
                    for (int m = kGroup / 2; m < kGroup; ++m) tc::umma_bf16_lohi(tmem + (m >> 2) * N % 256, a_lo + (m >> 2) * 1024 + (m & 3) * 2, b_lo + (m & 3) * 2, hi, idesc, 1);
                    tc::umma_commit(&bars[gi & 7]);
                }
                continue;
            }
            if (leader) {
#pragma unroll
                for (int m = 0; m < kGroup; ++m) tc::umma_bf16_lohi(tmem + (m >> 2) * N % 256, a_lo + (m >> 2) * 1024 + (m & 3) * 2, b_lo + (m & 3) * 2, hi, idesc, 1);
                if (kMode >= 1) tc::umma_commit(&bars[gi & 7]);
            }
        }
        if (leader) tc::umma_commit(&bar);
        tc::mbar_wait(&bar, 0);
        long long t1 = clock64();
        if (leader) out[blockIdx.x] = t1 - t0;
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc::tc_fence_after(); tc::tmem_dealloc<512>(tmem); }
}
template <int N, int G, int M>
void run_group(int grid) {
    long long* d; cudaMalloc(&d, sizeof(long long) * grid);
    const int groups = 2048;
    const size_t smem = 113 * 1024;
    cudaFuncSetAttribute(group_kernel<N, G, M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    group_kernel<N, G, M><<<grid, 128, smem>>>(d, 16);
    group_kernel<N, G, M><<<grid, 128, smem>>>(d, groups);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("N=%d groups of %d MMAs, mode %d (%s): %.1f cycles per group = %.1f per MMA   [%s]\n", N, G, M,
           M == 0 ? "no commit" : M == 1 ? "commit per group" : M == 2 ? "wait+fence before the group, commit" : "wait+fence in the middle of the group, commit", (double)mx / groups, (double)mx / groups / G, cudaGetErrorString(e));
    cudaFree(d);
}

template <int N, int D>
void run(const char* tag, int grid, int shift = 0) {
    long long* d; cudaMalloc(&d, sizeof(long long) * grid);
    const int iters = 4096;
    const size_t smem = 113 * 1024;
    cudaFuncSetAttribute(rate_kernel<N, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    rate_kernel<N, D><<<grid, 128, smem>>>(d, 64, shift);
    rate_kernel<N, D><<<grid, 128, smem>>>(d, iters, shift);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, d, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%s N=%d grid=%d A start +%d rows: %.1f cycles per MMA (M=128,K=16)  -> %.0f FLOP/clk/SM   [%s]\n", tag, N, grid, shift, (double)mx / (iters * 4.0),
           2.0 * 128 * N * 16 * iters * 4.0 / mx, cudaGetErrorString(e));
    cudaFree(d);
}
int main() {
    run_group<64, 8, 0>(148); run_group<64, 8, 1>(148); run_group<64, 8, 2>(148); run_group<64, 12, 1>(148); run_group<64, 12, 2>(148);
    run_group<128, 8, 1>(148); run_group<128, 8, 2>(148); run_group<64, 4, 2>(148);
    run_group<64, 4, 3>(148); run_group<64, 8, 3>(148); run_group<128, 4, 3>(148); run_group<128, 4, 2>(148);
    run<64, 0>("same A ", 148); run<128, 0>("same A ", 148); run<256, 0>("same A ", 148);
    run<64, 1>("4 A tiles", 148); run<128, 1>("4 A tiles", 148); run<256, 1>("4 A tiles", 148);
    for (int sh : {1, 2, 4, 7, 34, 35}) { run<64, 1>("4 A tiles", 148, sh); run<128, 1>("4 A tiles", 148, sh); }
    return 0;
}
