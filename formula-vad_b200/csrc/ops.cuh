// ops.cuh -- operator-level helpers (see ops.cu).
#pragma once
#include "common.cuh"

namespace fvad {

class FftPlan {  // device plan behind fvad_fft (reference src/FFT.zig)
public:
    explicit FftPlan(size_t n_fft);
    ~FftPlan();
    FftPlan(const FftPlan &) = delete;
    FftPlan &operator=(const FftPlan &) = delete;
    size_t n() const { return n_; }
    // device pointers; window may be nullptr
    void forward(const float *frames, size_t n_frames, size_t stride, const float *window, fvad_complex *bins, cudaStream_t s);
    void inverse(const fvad_complex *bins, size_t n_frames, float *out, cudaStream_t s);

private:
    struct Impl;
    size_t n_;
    Impl *impl_;
};

void launch_upsample(const float *y16, size_t ypitch, float *out, size_t opitch, size_t n16, int D, int rows, cudaStream_t s);

}  // namespace fvad
