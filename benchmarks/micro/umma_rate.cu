// Micro-benchmark: tcgen05.mma issue rate of ONE CTA per SM (M=128, N=256, K=16, bf16, operands in shared memory),
// alone (mode 0) and while a second thread streams 48 KB bulk copies into other shared-memory stages (mode 1).
// Answers: is the single-CTA GEMM limited by the tensor pipe, by shared-memory read bandwidth, or by smem fill traffic?
#include "../../csrc/common/picd_common.cuh"
#include <cstdio>
using namespace picd;

__global__ void __launch_bounds__(128, 1) umma_rate(const uint8_t* src, long long* cycles, int iters, int mode, int n) {
    extern __shared__ uint8_t raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    uint64_t* lbar = bar + 1;
    uint32_t* tptr = reinterpret_cast<uint32_t*>(smem + 64);
    uint8_t* sA = smem + 1024;
    uint8_t* sB = sA + 16384;
    uint8_t* sX = sB + 32768;      // 3 x 48 KB landing zone for the streaming copies
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(lbar, 1); fence_mbar_init(); }
    if (warp == 0) { tmem_alloc(tptr, 512); tmem_relinquish(); }
    for (int i = threadIdx.x; i < 49152 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sA)[i] = 0x3c003c00u;
    tc_fence_before(); __syncthreads(); tc_fence_after();
    fence_proxy_async_smem();
    const uint32_t tmem = *tptr;
    if (warp == 1 && lane == 0 && mode == 1) {
        uint32_t ph = 0;
        const uint8_t* g = src + (size_t)blockIdx.x * 49152 * 4;
        for (int it = 0; it < iters; ++it) {
            mbar_arrive_expect_tx(lbar, 49152);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(sX + (it % 3) * 49152)), "l"(reinterpret_cast<uint64_t>(g + (it & 3) * 49152)), "r"(49152), "r"(smem_u32(lbar)) : "memory");
            mbar_wait(lbar, ph); ph ^= 1;
        }
    }
    if (threadIdx.x == 0) {
        const uint32_t idesc = make_idesc_bf16(128, n, 0, 0);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                umma_bf16(tmem, make_smem_desc(smem_u32(sA) + k * 32, 16, 1024), make_smem_desc(smem_u32(sB) + k * 32, 16, 1024), idesc, 1);
        }
        umma_commit(bar);
        mbar_wait(bar, 0);
        const long long t1 = clock64();
        cycles[blockIdx.x] = t1 - t0;
    }
    tc_fence_before(); __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint8_t* src; cudaMalloc(&src, (size_t)sms * 49152 * 4); cudaMemset(src, 0, (size_t)sms * 49152 * 4);
    long long* cyc; cudaMallocManaged(&cyc, sms * sizeof(long long));
    cudaFuncSetAttribute(umma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const int iters = 2000;
    for (int n : {256, 128, 64}) for (int mode = 0; mode < 2; ++mode) {
        umma_rate<<<sms, 128, 200 * 1024>>>(src, cyc, iters, mode, n);
        cudaError_t e = cudaDeviceSynchronize();
        double avg = 0; for (int i = 0; i < sms; ++i) avg += cyc[i]; avg /= sms;
        printf("N=%3d mode=%d (%s): %s  %.1f cycles per MMA (K=16), %.0f MAC/clk/SM\n", n, mode, mode ? "MMA + 48KB bulk copy per 4 MMAs" : "MMA only",
               cudaGetErrorString(e), avg / (iters * 4.0), 128.0 * n * 16 * iters * 4 / avg);
    }
    return 0;
}
