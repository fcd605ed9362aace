// mmabw_pair2.cu -- does the way the GRU issues its MMAs cost tensor-pipe time?  N = 48 pair MMAs (cta_group::2, A from TMEM),
// groups of 25 accumulating MMAs (one GRU chunk) with a commit after each group:
//   A: one thread, one accumulator          B: one thread, six accumulators in rotation
//   C: two threads (warps), one accumulator each, issuing concurrently        D: two threads, three accumulators each
// Reports cycles per MMA (ideal 24).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../formula-vad_b200/csrc/tc_dev.cuh"
using namespace fvad::tc;

template <int NTHREADS_ISSUE, int NACC>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k(int groups, unsigned long long *cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar[8];
    __shared__ uint32_t tbase;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bar[i], 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc2<512>(&tbase);
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tmem = tbase;
    const int warp = threadIdx.x >> 5;
    if (warp < NTHREADS_ISSUE && cluster_ctarank() == 0) {
        constexpr uint32_t idesc = make_idesc_f16(256, 48);
        const long long t0 = clock64();
        uint32_t ph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int g = warp; g < groups; g += NTHREADS_ISSUE) {
            const int acc = (NTHREADS_ISSUE == 1) ? (g % NACC) : (warp + NTHREADS_ISSUE * ((g / NTHREADS_ISSUE) % NACC));
            const uint32_t d = tmem + 200 + 48 * acc;
            // B: rotate through 5 stages of 19 200 bytes like the GRU's ring
            const uint64_t bd0 = make_smem_desc(smem_u32(smem + (g % 5) * 19200), 24 * 16, 128);
            const uint32_t lo = (uint32_t)bd0, hi = (uint32_t)(bd0 >> 32);
            if (elect_one()) {
                if (g >= NTHREADS_ISSUE * NACC) {  // the accumulator's previous group has completed (like acc_empty)
                    mbar_wait(&bar[acc], ph[acc]);
                    ph[acc] ^= 1;
                    tc_fence_after();
                }
                mma2_ts_lohi<false>(d, tmem, lo, hi, idesc);
#pragma unroll
                for (int kk = 1; kk < 25; ++kk) mma2_ts_lohi<true>(d, tmem + kk * 8, lo + kk * 48, hi, idesc);
                mma2_commit_mc(&bar[acc], 1);
            }
            __syncwarp();
        }
        // drain
        if (elect_one())
            for (int a2 = 0; a2 < 8; ++a2) {
                bool mine = (NTHREADS_ISSUE == 1) ? (a2 < NACC) : (a2 % NTHREADS_ISSUE == warp && a2 < NTHREADS_ISSUE * NACC);
                if (mine && groups > a2) mbar_wait(&bar[a2], ph[a2]);
            }
        if ((threadIdx.x & 31) == 0) cycles[(blockIdx.x >> 1) * 2 + warp] = (unsigned long long)(clock64() - t0);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (threadIdx.x < 32) tmem_dealloc2<512>(tmem);
}

template <int NT, int NACC>
void run(unsigned long long *d, const char *name) {
    const int groups = 3000;
    cudaMemset(d, 0, 148 * 2 * 8);
    cudaFuncSetAttribute(k<NT, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    k<NT, NACC><<<148, 128, 96 * 1024>>>(groups, d);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 74; ++i) { double m = h[2 * i] > h[2 * i + 1] ? h[2 * i] : h[2 * i + 1]; avg += m; } avg /= 74;
    printf("%s: %s  %.1f cycles/MMA (ideal 24.0)\n", name, cudaGetErrorString(e), avg / (groups * 25.0));
}

int main() {
    unsigned long long *d; cudaMalloc(&d, 148 * 2 * 8);
    run<1, 1>(d, "A one issuer, 1 accumulator ");
    run<1, 2>(d, "B one issuer, 2 accumulators");
    run<1, 6>(d, "B one issuer, 6 accumulators");
    run<2, 1>(d, "C two issuers, 1 acc each   ");
    run<2, 3>(d, "D two issuers, 3 acc each   ");
    return 0;
}
