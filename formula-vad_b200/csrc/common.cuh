// common.cuh -- shared declarations of libfvad_cuda (host + device).
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

#include "../../include/fvad.h"

namespace fvad {

// ---- error plumbing: C++ exceptions stay inside the library; capi.cpp maps them to codes ----
struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

void set_last_error(const std::string &msg);
const std::string &last_error();

#define FVAD_CUDA_CHECK(expr)                                                                       \
    do {                                                                                            \
        cudaError_t e__ = (expr);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
            throw ::fvad::Error(FVAD_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__)); \
    } while (0)

// launch counter (fvad_batch_stats.kernel_launches; bench.py reports it as gpu_launches)
extern uint64_t g_kernel_launches;
#define FVAD_LAUNCHED() (++::fvad::g_kernel_launches)

// ---- NSNet2 / pipeline constants (src/NSNet2.zig:12-16) ----
constexpr int kNfft = FVAD_NS_NFFT;    // 320
constexpr int kHop = FVAD_NS_HOP;      // 160
constexpr int kBins = FVAD_NS_BINS;    // 161
constexpr int kFrames = FVAD_NS_FRAMES;// 50 frames per chunk
constexpr int kCtx = FVAD_NS_CTX;      // 4 context frames
constexpr int kSteps = FVAD_NS_STEPS;  // 54
constexpr int kChunk16 = FVAD_NS_CHUNK16; // 8000
constexpr int kYLook = 1024;           // y16 samples kept in front of a span for VAD re-blocking

constexpr int kHidden = 400;
constexpr int kGates = 1200;

template <typename T>
struct DevBuf {  // owning device buffer, grow-only
    T *p = nullptr;
    size_t n = 0;
    void ensure(size_t count) {
        if (count <= n) return;
        if (p) cudaFree(p);
        p = nullptr;
        cudaError_t e = cudaMalloc(&p, count * sizeof(T));
        if (e != cudaSuccess) {
            n = 0;
            throw Error(FVAD_ERR_OUT_OF_MEMORY, std::string("cudaMalloc: ") + cudaGetErrorString(e));
        }
        n = count;
    }
    void zero(cudaStream_t s) { if (p) cudaMemsetAsync(p, 0, n * sizeof(T), s); }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
    ~DevBuf() { release(); }
    DevBuf() = default;
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
};

}  // namespace fvad
