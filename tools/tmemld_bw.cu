// tmemld_bw.cu -- tcgen05.ld throughput per SM: alone and next to a saturated tcgen05.mma stream (A from TMEM, N = 144).
// The question it answers: how fast can an epilogue drain fp32 accumulators (K = 400 GEMMs need 41 B/clk/SM at the full
// tensor rate) and does the drain slow the MMAs down.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../formula-vad_b200/csrc/tc_dev.cuh"
using namespace fvad::tc;

__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&r)[64]) {
    uint32_t a[32], b[32];
    tmem_ld32(taddr, a); tmem_ld32(taddr + 32, b);
#pragma unroll
    for (int i = 0; i < 32; ++i) { r[i] = a[i]; r[32 + i] = b[i]; }
}

// other shapes, 32 registers per thread each: 16x256b.x8 (16 lanes x 8 x 256 bit, two of them cover a warp's 32 lanes),
// 16x128b.x16, and the 32x32b shape as one x64 / x128 instruction
#define LD_REGS32(r) "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), \
    "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), \
    "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
#define LD_FMT32 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
__device__ __forceinline__ void ld_16x256b_x8(uint32_t t, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 " LD_FMT32 : LD_REGS32(r) : "r"(t) : "memory");
}
__device__ __forceinline__ void ld_16x128b_x16(uint32_t t, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.16x128b.x16.b32 " LD_FMT32 : LD_REGS32(r) : "r"(t) : "memory");
}

template <int SHAPE>
__global__ void __launch_bounds__(128, 1) shape_kernel(int reps, unsigned long long *out) {
    __shared__ uint32_t tbase;
    if (threadIdx.x < 32) tmem_alloc<512>(&tbase);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tbase;
    const int warp = threadIdx.x >> 5;
    const uint32_t t = tmem + ((uint32_t)(warp * 32) << 16);
    unsigned sink = 0;
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
        uint32_t a[32], b[32];
        if (SHAPE == 0) { ld_16x256b_x8(t, a); ld_16x256b_x8(t + (16u << 16), b); }        // 2 x (16 lanes x 64 columns) = 8 KB
        if (SHAPE == 1) { ld_16x128b_x16(t, a); ld_16x128b_x16(t + (16u << 16), b); }      // 2 x (16 lanes x 64 columns) = 8 KB
        if (SHAPE == 2) { tmem_ld32(t, a); tmem_ld32(t + 32, b); }                          // 2 x (32 lanes x 32 columns) = 8 KB
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) sink ^= a[i] ^ b[i];
    }
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) out[blockIdx.x * 16 + warp] = (unsigned long long)(t1 - t0);
    if (sink == 0x12345678u) out[0] = sink;
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem);
}
template <int SHAPE>
void run_shape(unsigned long long *d, const char *name) {
    const int reps = 4000;
    cudaMemset(d, 0, 148 * 16 * 8);
    shape_kernel<SHAPE><<<148, 128>>>(reps, d);
    cudaError_t e = cudaDeviceSynchronize();
    static unsigned long long h[148 * 16];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = 0;
    for (int b = 0; b < 148; ++b) { double mx = 0; for (int w = 0; w < 4; ++w) mx = mx > (double)h[b * 16 + w] ? mx : (double)h[b * 16 + w]; cyc += mx; }
    cyc /= 148;
    printf("shape %s, 4 warps x 8 KB per iteration: %s  %.1f B/clk/SM\n", name, cudaGetErrorString(e), (double)reps * 4 * 8192 / cyc);
}

// LDW warps (4 or 8; warp w reads lane quadrant w % 4) each issue `reps` x (DEPTH loads of 32 columns, then one wait).
// With MMA = true one elected thread of an extra warp keeps the tensor pipe busy with 128 x 144 x 16 MMAs.
template <int LDW, int DEPTH, bool MMA>
__global__ void __launch_bounds__(32 * (LDW + 1), 1) ld_kernel(int reps, unsigned long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tbase;
    __shared__ volatile int stop;
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); stop = 0; }
    if (threadIdx.x < 32) tmem_alloc<512>(&tbase);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tbase;
    const int warp = threadIdx.x >> 5;
    unsigned sink = 0;
    if (warp < LDW) {
        const uint32_t t = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (warp >= 4 ? 128u : 0u);
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r) {
            uint32_t v[DEPTH][32];
#pragma unroll
            for (int d = 0; d < DEPTH; ++d) tmem_ld32(t + 32 * (d & 3), v[d]);
            tmem_wait_ld();
#pragma unroll
            for (int d = 0; d < DEPTH; ++d)
#pragma unroll
                for (int i = 0; i < 32; ++i) sink ^= v[d][i];
        }
        const long long t1 = clock64();
        if ((threadIdx.x & 31) == 0) out[blockIdx.x * 16 + warp] = (unsigned long long)(t1 - t0);
        __syncwarp();
        if (threadIdx.x == 0) stop = 1;
    } else if (MMA) {
        if (elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(128, 144);
            const uint64_t bd0 = make_smem_desc(smem_u32(smem), 144 * 16, 128);
            const uint32_t lo = (uint32_t)bd0, hi = (uint32_t)(bd0 >> 32);
            unsigned long long n = 0;
            const long long t0 = clock64();
            while (!stop) {
#pragma unroll
                for (int kk = 0; kk < 25; ++kk) {
                    if (kk) mma_ts_lohi<true>(tmem + 256, tmem + 400 + (kk & 7) * 8, lo + (kk & 3) * (2 * 144), hi, idesc);
                    else mma_ts_lohi<false>(tmem + 256, tmem + 400, lo, hi, idesc);
                }
                mma_commit(&bar);
                mbar_wait(&bar, n & 1);
                ++n;
            }
            out[blockIdx.x * 16 + 8] = (unsigned long long)(clock64() - t0);
            out[blockIdx.x * 16 + 9] = n * 25;
        }
    }
    if (sink == 0x12345678u) out[0] = sink;
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem);
}

template <int LDW, int DEPTH, bool MMA>
void run(unsigned long long *d) {
    const int reps = 4000;
    cudaMemset(d, 0, 148 * 16 * 8);
    cudaFuncSetAttribute(ld_kernel<LDW, DEPTH, MMA>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    ld_kernel<LDW, DEPTH, MMA><<<148, 32 * (LDW + 1), 64 * 1024>>>(reps, d);
    cudaError_t e = cudaDeviceSynchronize();
    static unsigned long long h[148 * 16];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    double cyc = 0, mcyc = 0, mn = 0;
    for (int b = 0; b < 148; ++b) {
        double mx = 0;
        for (int w = 0; w < LDW; ++w) mx = mx > (double)h[b * 16 + w] ? mx : (double)h[b * 16 + w];
        cyc += mx; mcyc += (double)h[b * 16 + 8]; mn += (double)h[b * 16 + 9];
    }
    cyc /= 148;
    const double bytes = (double)reps * DEPTH * LDW * 32 * 32 * 4;
    printf("ld warps %d depth %d %s: %s  %.1f B/clk/SM  (%.0f cycles per 4 KB warp-load)", LDW, DEPTH, MMA ? "+MMA" : "alone", cudaGetErrorString(e),
           bytes / cyc, cyc / ((double)reps * DEPTH));
    if (MMA && mn > 0) printf("   MMA: %.1f cycles per 128x144x16 (ideal 72)", mcyc / mn);
    printf("\n");
}

int main() {
    unsigned long long *d; cudaMalloc(&d, 148 * 16 * 8);
    run_shape<2>(d, "32x32b.x32 x2"); run_shape<0>(d, "16x256b.x8 x2"); run_shape<1>(d, "16x128b.x16 x2");
    run<4, 1, false>(d); run<4, 2, false>(d); run<4, 4, false>(d); run<8, 1, false>(d); run<8, 2, false>(d);
    run<4, 1, true>(d); run<4, 2, true>(d); run<4, 4, true>(d); run<8, 2, true>(d);
    return 0;
}
