// mmabw.cu -- microbenchmark: tcgen05.mma issue rate vs N (M=128, K=16 per instruction, kind::f16),
// A from shared memory (SS) or from tensor memory (TS).  One CTA per SM, one issuing thread.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../formula-vad_b200/csrc/tc_dev.cuh"
using namespace fvad::tc;

constexpr int kMmas = 4000;

__global__ void __launch_bounds__(128, 1) mma_kernel(int N, int a_tmem, int kgroup, unsigned long long *cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc<512>(&tbase);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tbase;
    if (threadIdx.x == 0) {
        const uint32_t idesc = make_idesc_f16(128, N);
        const uint32_t sa = smem_u32(smem), sb = smem_u32(smem + 110 * 1024);
        const long long t0 = clock64();
        for (int i = 0, kk = 0; i < kMmas; ++i, kk = (kk + 1 == kgroup) ? 0 : kk + 1) {  // walk K like the real kernels: kgroup K-steps accumulate into one tile
            const uint64_t bd = make_smem_desc(sb + (kk & 3) * 2 * (N * 16), N * 16, 128);  // stays inside 25 KB
            if (a_tmem) mma_ts(tmem + 256, tmem + kk * 8, bd, idesc, kk > 0);
            else mma_ss(tmem + 256, make_smem_desc(sa + kk * 2 * (128 * 16), 128 * 16, 128), bd, idesc, kk > 0);
        }
        mma_commit(&bar);
        mbar_wait(&bar, 0);
        cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem);
}

int main() {
    unsigned long long *d; cudaMalloc(&d, 148 * 8);
    cudaFuncSetAttribute(mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int Ns[] = {16, 32, 48, 64, 96, 128, 192, 240, 256};
    for (int a_tmem = 0; a_tmem < 2; ++a_tmem)
        for (int N : Ns) {
            mma_kernel<<<148, 128, 200 * 1024>>>(N, a_tmem, 25, d);
            cudaError_t e = cudaDeviceSynchronize();
            unsigned long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)h[i]; avg /= 148;
            printf("A=%s N=%3d: %s  %.1f cycles/MMA  (ideal %.1f)  efficiency %.0f%%\n", a_tmem ? "tmem" : "smem", N,
                   cudaGetErrorString(e), avg / kMmas, N / 2.0, 100.0 * (N / 2.0) / (avg / kMmas));
        }
    return 0;
}
