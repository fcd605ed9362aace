// mmabw_pair.cu -- tcgen05.mma.cta_group::2 (M = 256 over a CTA pair, A from TMEM, B halves in both CTAs' shared memory):
// cycles per MMA for a fully unrolled single-thread issue loop, per N.  Ideal = N/2 cycles (same as one CTA's M = 128).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../formula-vad_b200/csrc/tc_dev.cuh"
using namespace fvad::tc;

template <int N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) mma2_kernel(int reps, unsigned long long *cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc2<512>(&tbase);
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tmem = tbase;
    if (threadIdx.x < 32 && cluster_ctarank() == 0) {
        constexpr uint32_t idesc = make_idesc_f16(256, N);
        const uint64_t bd0 = make_smem_desc(smem_u32(smem), (N / 2) * 16, 128);
        const uint32_t lo = (uint32_t)bd0, hi = (uint32_t)(bd0 >> 32);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < 25; ++kk) {
                    if (kk) mma2_ts_lohi<true>(tmem + 256, tmem + kk * 8, lo + (kk & 3) * N, hi, idesc);
                    else mma2_ts_lohi<false>(tmem + 256, tmem, lo, hi, idesc);
                }
            }
            __syncwarp();
        }
        if (elect_one()) mma2_commit_mc(&bar, 1);
        mbar_wait(&bar, 0);
        if (threadIdx.x == 0) cycles[blockIdx.x >> 1] = (unsigned long long)(clock64() - t0);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (threadIdx.x < 32) tmem_dealloc2<512>(tmem);
}

template <int N>
void run(unsigned long long *d) {
    const int reps = 200;
    cudaFuncSetAttribute(mma2_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    mma2_kernel<N><<<148, 128, 64 * 1024>>>(reps, d);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[74]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 74; ++i) avg += (double)h[i]; avg /= 74;
    const double per = avg / (reps * 25.0);
    printf("pair TS unrolled N=%3d: %s  %.1f cycles/MMA  (ideal %.1f)  efficiency %.0f%%\n", N, cudaGetErrorString(e), per, N / 2.0, 100.0 * (N / 2.0) / per);
}

int main() {
    unsigned long long *d; cudaMalloc(&d, 148 * 8);
    run<48>(d); run<64>(d); run<96>(d); run<144>(d); run<208>(d); run<256>(d);
    return 0;
}
