// mmabw2.cu -- tcgen05.mma (A from TMEM) issue floor with a fully unrolled, immediate-offset issue loop.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../formula-vad_b200/csrc/tc_dev.cuh"
using namespace fvad::tc;

__device__ __forceinline__ void mma_ts_lohi(uint32_t d, uint32_t a, uint32_t blo, uint32_t bhi, uint32_t idesc, bool acc) {
    if (acc)
        asm volatile("{\n\t.reg .b64 bd;\n\t.reg .pred p;\n\tsetp.eq.u32 p, 1, 1;\n\tmov.b64 bd, {%2, %3};\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n\t}\n" ::"r"(d), "r"(a), "r"(blo), "r"(bhi), "r"(idesc) : "memory");
    else
        asm volatile("{\n\t.reg .b64 bd;\n\t.reg .pred p;\n\tsetp.eq.u32 p, 1, 0;\n\tmov.b64 bd, {%2, %3};\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n\t}\n" ::"r"(d), "r"(a), "r"(blo), "r"(bhi), "r"(idesc) : "memory");
}

__device__ __forceinline__ void mma_ss_lohi(uint32_t d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc, bool acc) {
    if (acc)
        asm volatile("{\n\t.reg .b64 ad, bd;\n\t.reg .pred p;\n\tsetp.eq.u32 p, 1, 1;\n\tmov.b64 ad, {%1, %2};\n\tmov.b64 bd, {%3, %4};\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, p;\n\t}\n" ::"r"(d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc) : "memory");
    else
        asm volatile("{\n\t.reg .b64 ad, bd;\n\t.reg .pred p;\n\tsetp.eq.u32 p, 1, 0;\n\tmov.b64 ad, {%1, %2};\n\tmov.b64 bd, {%3, %4};\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %5, p;\n\t}\n" ::"r"(d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc) : "memory");
}

template <int N, bool SS>
__global__ void __launch_bounds__(128, 1) mma_kernel(int reps, unsigned long long *cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < 150 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc<512>(&tbase);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tbase;
    if (threadIdx.x < 32) {
        constexpr uint32_t idesc = make_idesc_f16(128, N);
        const uint64_t bd0 = make_smem_desc(smem_u32(smem), N * 16, 128);
        const uint32_t lo = (uint32_t)bd0, hi = (uint32_t)(bd0 >> 32);
        const uint64_t ad0 = make_smem_desc(smem_u32(smem + 40 * 1024), 128 * 16, 128);
        const uint32_t alo = (uint32_t)ad0, ahi = (uint32_t)(ad0 >> 32);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < 25; ++kk)
                    if (SS) mma_ss_lohi(tmem + 256, alo + kk * 256, ahi, lo + (kk & 3) * (2 * N), hi, idesc, kk > 0);
                    else mma_ts_lohi(tmem + 256, tmem + kk * 8, lo + (kk & 3) * (2 * N), hi, idesc, kk > 0);
            }
            __syncwarp();
        }
        if (elect_one()) { mma_commit(&bar); }
        mbar_wait(&bar, 0);
        if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem);
}

template <int N, bool SS>
void run(unsigned long long *d) {
    const int reps = 200;
    cudaFuncSetAttribute(mma_kernel<N, SS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 150 * 1024);
    mma_kernel<N, SS><<<148, 128, 150 * 1024>>>(reps, d);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)h[i]; avg /= 148;
    const double per = avg / (reps * 25.0);
    printf("%s unrolled N=%3d: %s  %.1f cycles/MMA  (ideal %.1f)  efficiency %.0f%%\n", SS ? "SS" : "TS", N, cudaGetErrorString(e), per, N / 2.0, 100.0 * (N / 2.0) / per);
}

int main() {
    unsigned long long *d; cudaMalloc(&d, 148 * 8);
    run<48, false>(d); run<240, false>(d); run<48, true>(d); run<96, true>(d); run<128, true>(d); run<192, true>(d); run<208, true>(d); run<240, true>(d); run<256, true>(d);
    return 0;
}
