// ops.cu -- stand-alone operator kernels behind the reference's FFT.zig call surface
// (FFT.fft / FFT.invFft, src/FFT.zig:85-134) and the 1:D upsampler that turns the 16 kHz denoised
// stream back into input-rate PCM (audio_utils/resample.zig:32-79; NSNet2.zig:231-236).
#include <cmath>
#include <vector>

#include "fft_dev.cuh"
#include "ops.cuh"

namespace fvad {

// One warp per frame, real length n = 64*R.  T: exp(-2*pi*i*j/n), j<n (global; staged in smem).
template <int R>
__global__ void __launch_bounds__(128) rfft_fwd_kernel(const float *__restrict__ frames, size_t n_frames, size_t stride,
                                                       const float *__restrict__ window, float2 *__restrict__ bins,
                                                       const float2 *__restrict__ gT) {
    constexpr int N = 32 * R, NR = 2 * N;
    extern __shared__ float2 sm[];
    float2 *T = sm;                       // NR entries
    float2 *Zall = sm + NR;               // 4 warps x N
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < NR; i += 128) T[i] = gT[i];
    __syncthreads();
    float2 *Z = Zall + warp * N;
    for (size_t f = (size_t)blockIdx.x * 4 + warp; f < n_frames; f += (size_t)gridDim.x * 4) {
        const float *x = frames + f * stride;
        float2 v[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
            const int n = lane + 32 * q;
            float a = x[2 * n], b = x[2 * n + 1];
            if (window) { a *= window[2 * n]; b *= window[2 * n + 1]; }  // FFT.zig:188-196
            v[q] = make_float2(a, b);
        }
        warp_cfft<R, false>(v, T, lane);
#pragma unroll
        for (int k1 = 0; k1 < R; ++k1) Z[cfft_out_index<R>(lane, k1)] = v[k1];
        __syncwarp();
        for (int k = lane; k <= N; k += 32) bins[f * (N + 1) + k] = rfft_bin(Z, T, N, k);
        __syncwarp();
    }
}

template <int R>
__global__ void __launch_bounds__(128) rfft_inv_kernel(const float2 *__restrict__ bins, size_t n_frames,
                                                       float *__restrict__ out, const float2 *__restrict__ gT) {
    constexpr int N = 32 * R, NR = 2 * N;
    extern __shared__ float2 sm[];
    float2 *T = sm;
    float2 *Xall = sm + NR;               // 4 warps x (N+1)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < NR; i += 128) T[i] = gT[i];
    __syncthreads();
    float2 *X = Xall + warp * (N + 1);
    for (size_t f = (size_t)blockIdx.x * 4 + warp; f < n_frames; f += (size_t)gridDim.x * 4) {
        for (int k = lane; k <= N; k += 32) X[k] = bins[f * (N + 1) + k];
        __syncwarp();
        float2 v[R];
#pragma unroll
        for (int q = 0; q < R; ++q) v[q] = irfft_pack(X, T, N, lane + 32 * q);
        warp_cfft<R, true>(v, T, lane);
        float2 *o = reinterpret_cast<float2 *>(out + f * NR);
#pragma unroll
        for (int k1 = 0; k1 < R; ++k1) o[cfft_out_index<R>(lane, k1)] = v[k1];
        __syncwarp();
    }
}

struct FftPlan::Impl {
    int R = 0;
    float2 *T = nullptr;
};

FftPlan::FftPlan(size_t n_fft) : n_(n_fft), impl_(new Impl) {
    // FFT.init, src/FFT.zig:41-43
    if (n_fft == 0 || (n_fft & 1)) { delete impl_; throw Error(FVAD_ERR_INVALID_FFT_SIZE, "n_fft must be even and non-zero"); }
    const int R = (int)(n_fft / 64);
    if (n_fft % 64 || !(R == 1 || R == 2 || R == 4 || R == 5 || R == 8 || R == 16)) {
        delete impl_;
        throw Error(FVAD_ERR_INVALID_FFT_SIZE, "GPU FFT sizes: 64, 128, 256, 320, 512, 1024");
    }
    impl_->R = R;
    std::vector<float2> t(n_fft);
    const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
    for (size_t j = 0; j < n_fft; ++j) {
        const double ph = -2.0 * pi * (double)j / (double)n_fft;
        t[j] = make_float2((float)std::cos(ph), (float)std::sin(ph));
    }
    if (cudaMalloc(&impl_->T, n_fft * sizeof(float2)) != cudaSuccess) { delete impl_; throw Error(FVAD_ERR_OUT_OF_MEMORY, "cudaMalloc"); }
    cudaMemcpy(impl_->T, t.data(), n_fft * sizeof(float2), cudaMemcpyHostToDevice);
}

FftPlan::~FftPlan() {
    if (impl_) { cudaFree(impl_->T); delete impl_; }
}

template <int R>
static void fwd(const float *frames, size_t nf, size_t stride, const float *win, float2 *bins, const float2 *T, cudaStream_t s) {
    constexpr int N = 32 * R;
    const size_t smem = (2 * N + 4 * N) * sizeof(float2);
    const unsigned grid = (unsigned)std::min<size_t>((nf + 3) / 4, 148 * 16);
    rfft_fwd_kernel<R><<<grid, 128, smem, s>>>(frames, nf, stride, win, bins, T);
}
template <int R>
static void inv(const float2 *bins, size_t nf, float *out, const float2 *T, cudaStream_t s) {
    constexpr int N = 32 * R;
    const size_t smem = (2 * N + 4 * (N + 1)) * sizeof(float2);
    const unsigned grid = (unsigned)std::min<size_t>((nf + 3) / 4, 148 * 16);
    rfft_inv_kernel<R><<<grid, 128, smem, s>>>(bins, nf, out, T);
}

void FftPlan::forward(const float *frames, size_t nf, size_t stride, const float *win, fvad_complex *bins, cudaStream_t s) {
    if (!nf) return;
    float2 *b = reinterpret_cast<float2 *>(bins);
    switch (impl_->R) {
        case 1: fwd<1>(frames, nf, stride, win, b, impl_->T, s); break;
        case 2: fwd<2>(frames, nf, stride, win, b, impl_->T, s); break;
        case 4: fwd<4>(frames, nf, stride, win, b, impl_->T, s); break;
        case 5: fwd<5>(frames, nf, stride, win, b, impl_->T, s); break;
        case 8: fwd<8>(frames, nf, stride, win, b, impl_->T, s); break;
        case 16: fwd<16>(frames, nf, stride, win, b, impl_->T, s); break;
    }
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

void FftPlan::inverse(const fvad_complex *bins, size_t nf, float *out, cudaStream_t s) {
    if (!nf) return;
    const float2 *b = reinterpret_cast<const float2 *>(bins);
    switch (impl_->R) {
        case 1: inv<1>(b, nf, out, impl_->T, s); break;
        case 2: inv<2>(b, nf, out, impl_->T, s); break;
        case 4: inv<4>(b, nf, out, impl_->T, s); break;
        case 5: inv<5>(b, nf, out, impl_->T, s); break;
        case 8: inv<8>(b, nf, out, impl_->T, s); break;
        case 16: inv<16>(b, nf, out, impl_->T, s); break;
    }
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

// resample.zig:32-79: out[D*m + D-1] = in[m]; the D-1 samples before it interpolate from in[m-1]
// with std.math.lerp (a fused a + (b-a)*t), t = (i+1)/D.  `in` has one look-back sample at in[-1].
__global__ void upsample_kernel(const float *__restrict__ y16, size_t ypitch, float *__restrict__ out, size_t opitch,
                                size_t n16, int D, int rows) {
    const size_t total = n16 * (size_t)rows;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / n16, m = i % n16;
        const float *y = y16 + r * ypitch;
        const float cur = y[m], prev = y[(long long)m - 1];
        float *o = out + r * opitch + m * D;
        for (int j = 0; j < D - 1; ++j) o[j] = fmaf(cur - prev, __fdiv_rn((float)(j + 1), (float)D), prev);
        o[D - 1] = cur;
    }
}

void launch_upsample(const float *y16, size_t ypitch, float *out, size_t opitch, size_t n16, int D, int rows, cudaStream_t s) {
    if (!n16 || !rows) return;
    const size_t total = n16 * (size_t)rows;
    upsample_kernel<<<(unsigned)std::min<size_t>((total + 255) / 256, 148 * 16), 256, 0, s>>>(y16, ypitch, out, opitch, n16, D, rows);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
