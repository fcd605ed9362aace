// l2bw.cu -- microbenchmark: per-SM L2 -> shared-memory ingest rate of cp.async.bulk, plain vs
// cluster-of-2 multicast (each CTA fetches half of every stage and multicasts it to both).
// Decides whether the NSNet2 kernels' ~30 B/clk/SM operand stream is an SM-ingest or an L2-read limit.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../formula-vad_b200/csrc/tc_dev.cuh"
using namespace fvad::tc;

constexpr int kStages = 5, kStageBytes = 38400, kIters = 2000;  // 25 stages = one 960 KB weight image

template <int CL>
__global__ void __launch_bounds__(128, 1) bw_kernel(const uint8_t *src, unsigned long long *cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes);
    uint64_t *empty = full + kStages;
    const uint32_t rank = CL > 1 ? cluster_ctarank() : 0;
    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], CL); }
        fence_mbar_init();
    }
    if (CL > 1) cluster_sync(); else __syncthreads();
    const long long t0 = clock64();
    if (threadIdx.x == 0) {  // producer
        for (int it = 0; it < kIters; ++it) {
            const int st = it % kStages;
            if (CL > 1) mbar_wait_cluster(&empty[st], ((it / kStages) & 1) ^ 1); else mbar_wait(&empty[st], ((it / kStages) & 1) ^ 1);
            mbar_expect_tx(&full[st], kStageBytes);
            const uint8_t *s = src + (size_t)(it % 25) * kStageBytes;
            if (CL > 1) {
                const uint32_t half = kStageBytes / CL;
                bulk_g2s_mc(smem + st * kStageBytes + rank * half, s + rank * half, half, &full[st], (uint16_t)((1u << CL) - 1));
            } else {
                bulk_g2s(smem + st * kStageBytes, s, kStageBytes, &full[st]);
            }
        }
    } else if (threadIdx.x == 32) {  // consumer: frees the stage in every CTA of the cluster
        for (int it = 0; it < kIters; ++it) {
            const int st = it % kStages;
            mbar_wait(&full[st], (it / kStages) & 1);
            if (CL > 1) { for (uint32_t c = 0; c < CL; ++c) mbar_arrive_cluster(&empty[st], c); }
            else mbar_arrive(&empty[st]);
        }
    }
    __syncthreads();
    if (CL > 1) cluster_sync();
    if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
}

template <int CL>
void run(const uint8_t *src, unsigned long long *d_cyc, int grid) {
    const int smem = kStages * kStageBytes + 256;
    cudaFuncSetAttribute(bw_kernel<CL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        cudaLaunchKernelEx(&cfg, bw_kernel<CL>, src, d_cyc);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        unsigned long long h[160]; cudaMemcpy(h, d_cyc, grid * sizeof(h[0]), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid;
        const double bytes = (double)kIters * kStageBytes;
        printf("cluster=%d grid=%d: %s  %.3f ms  %.1f B/clk/SM ingested  chip %.2f TB/s delivered\n", CL, grid,
               cudaGetErrorString(e), ms, bytes / avg, bytes * grid / (ms * 1e-3) / 1e12);
    }
}

int main() {
    uint8_t *src; cudaMalloc(&src, 25 * kStageBytes); cudaMemset(src, 1, 25 * kStageBytes);
    unsigned long long *d_cyc; cudaMalloc(&d_cyc, 160 * sizeof(unsigned long long));
    run<1>(src, d_cyc, 148);
    run<1>(src, d_cyc, 74);
    run<2>(src, d_cyc, 148);
    run<4>(src, d_cyc, 148);
    return 0;
}
