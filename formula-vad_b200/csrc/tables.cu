// tables.cu -- constant tables (twiddles, windows) computed on the host with the reference's
// own recipes and uploaded once per device.
#include <cmath>
#include <vector>

#include "kernels.cuh"

namespace fvad {

uint64_t g_kernel_launches = 0;

static float2 *upload2(const std::vector<float2> &v) {
    float2 *d = nullptr;
    FVAD_CUDA_CHECK(cudaMalloc(&d, v.size() * sizeof(float2)));
    FVAD_CUDA_CHECK(cudaMemcpy(d, v.data(), v.size() * sizeof(float2), cudaMemcpyHostToDevice));
    return d;
}
static float *upload1(const std::vector<float> &v) {
    float *d = nullptr;
    FVAD_CUDA_CHECK(cudaMalloc(&d, v.size() * sizeof(float)));
    FVAD_CUDA_CHECK(cudaMemcpy(d, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
    return d;
}

// exp(-2*pi*i*j/n): phase in double, rounded to float (kiss_fft.c:361-367)
static std::vector<float2> twiddles(int n) {
    std::vector<float2> t((size_t)n);
    const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
    for (int j = 0; j < n; ++j) {
        const double ph = -2.0 * pi * j / n;
        t[(size_t)j] = make_float2((float)std::cos(ph), (float)std::sin(ph));
    }
    return t;
}

void Tables::build() {
    T320 = upload2(twiddles(320));
    T1024 = upload2(twiddles(1024));
    // sqrt(symmetric Hann(320)) in f32: window_fn.zig:30-41 + NSNet2.zig:384-396
    std::vector<float> w(320), wn(320);
    {
        const float N = 320.0f;
        const float step = (float)(2 * 3.14159265358979323846) / (N - 1);
        for (int i = 0; i < 320; ++i) {
            const float h = 0.5f - 0.5f * cosf((float)i * step);
            w[(size_t)i] = sqrtf(h);
            wn[(size_t)i] = w[(size_t)i] * (1.0f / 320.0f);  // window[i] * vol_norm_factor, NSNet2.zig:323,335
        }
    }
    win320 = upload1(w);
    win320n = upload1(wn);
    // periodic Hann(1024): window_fn.zig:22-28,51-68 (cosine-sum form in f32)
    std::vector<float> h(1024);
    {
        const float N = 1024.0f;
        const float two_pi = (float)(2 * 3.14159265358979323846);
        float sum = 0.0f;
        for (int i = 0; i < 1024; ++i) {
            float acc = 0.0f;
            acc += 1.0f * 0.5f * cosf((two_pi * 0.0f * (float)i) / N);
            acc += -1.0f * (1.0f - 0.5f) * cosf((two_pi * 1.0f * (float)i) / N);
            h[(size_t)i] = acc;
            sum += acc;  // windowNormFactor: sequential f32 sum, window_fn.zig:8-16
        }
        vad_norm = (1024.0f / sum) / 1024.0f;  // BufferedFFT.zig:99
    }
    hann1024 = upload1(h);
}

void Tables::release() {
    cudaFree(T320); cudaFree(T1024); cudaFree(win320); cudaFree(win320n); cudaFree(hann1024);
    T320 = T1024 = nullptr; win320 = win320n = hann1024 = nullptr;
}

__global__ void copy_rows_kernel(float *dst, const float *src, size_t dst_pitch, size_t src_pitch, size_t row_len, size_t rows) {
    const size_t total = row_len * rows;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / row_len, c = i % row_len;
        dst[r * dst_pitch + c] = src[r * src_pitch + c];
    }
}

void launch_copy_rows(float *dst, const float *src, size_t dst_pitch, size_t src_pitch, size_t row_len, size_t rows, cudaStream_t s) {
    if (!rows || !row_len) return;
    const size_t total = row_len * rows;
    const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 8);
    copy_rows_kernel<<<grid, 256, 0, s>>>(dst, src, dst_pitch, src_pitch, row_len, rows);
    FVAD_LAUNCHED();
}

// 16-bit PCM -> f32, x * 2^-15 (libsndfile's normalised float read, the reference's ingest:
// src/audio_utils/AudioFileStream.zig:56-102).  rows x n elements, 8 samples per thread.
__global__ void i16_to_f32_kernel(float *dst, const int16_t *src, size_t dst_pitch, size_t src_pitch, size_t n, size_t rows) {
    const size_t n8 = n / 8, per_row = n8 + 1;  // the last slot of a row handles the n % 8 tail
    const size_t total = per_row * rows;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / per_row, c = i % per_row;
        const int16_t *sp = src + r * src_pitch;
        float *dp = dst + r * dst_pitch;
        const bool vec = ((reinterpret_cast<uintptr_t>(sp) | reinterpret_cast<uintptr_t>(dp)) & 15) == 0;
        if (c < n8 && vec) {
            const uint4 v = *reinterpret_cast<const uint4 *>(sp + c * 8);
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
            float o[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                o[2 * k] = (float)(int16_t)(w[k] & 0xFFFFu) * (1.0f / 32768.0f);
                o[2 * k + 1] = (float)(int16_t)(w[k] >> 16) * (1.0f / 32768.0f);
            }
            *reinterpret_cast<float4 *>(dp + c * 8) = make_float4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<float4 *>(dp + c * 8 + 4) = make_float4(o[4], o[5], o[6], o[7]);
        } else if (c < n8) {
            for (size_t k = c * 8; k < c * 8 + 8; ++k) dp[k] = (float)sp[k] * (1.0f / 32768.0f);
        } else {
            for (size_t k = n8 * 8; k < n; ++k) dp[k] = (float)sp[k] * (1.0f / 32768.0f);
        }
    }
}

// pitches must be multiples of 8 elements and the bases 16-byte aligned (stage buffers are)
void launch_i16_to_f32(float *dst, const int16_t *src, size_t dst_pitch, size_t src_pitch, size_t n, size_t rows, cudaStream_t s) {
    if (!rows || !n) return;
    const size_t total = (n / 8 + 1) * rows;
    const int grid = (int)std::min<size_t>((total + 255) / 256, 148 * 16);
    i16_to_f32_kernel<<<grid, 256, 0, s>>>(dst, src, dst_pitch, src_pitch, n, rows);
    FVAD_LAUNCHED();
}

}  // namespace fvad
