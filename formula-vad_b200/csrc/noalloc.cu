// noalloc.cu -- fvad_vad_machine_scan_noalloc: the VAD machine operator (fvad_vad_machine_scan) for callers that cannot
// allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  Same kernels as the existing operator
// (vadscan.cu: RollingAverage x3 + the VADMachine state machine, src/AudioPipeline/VADMachine.zig:138-325), but the
// operator's scratch -- the derived machine table, the per-machine scan state and the three rolling-average rings per
// machine -- lives in ONE file-scope `static __device__` array of a fixed maximum size instead of cudaMalloc'd buffers.
// Every GPU that loads the library holds the array (~2.1 MB) until the process exits.  The caller passes inputs and
// outputs as DEVICE pointers.  This entry point calls no cudaMalloc / cudaFree and no allocating library routine.
//
// The array is shared by every call and keeps the previous call's contents: each call re-initialises what must start
// clean (state, rings, machine table) with a kernel of its own before the scan.  Calls that overlap on the GPU (two
// streams, two host threads) RACE on it: serialise them.
#include <climits>
#include <cstring>
#include <mutex>
#include <string>

#include "engine.cuh"
#include "kernels.cuh"

namespace fvad {
namespace {

constexpr int kMaxMachines = FVAD_NOALLOC_MAX_MACHINES;
constexpr int kMaxRing = FVAD_NOALLOC_MAX_RING;  // doubles per machine: long + short + ratio windows

struct Scratch {
    double rings[(size_t)kMaxMachines * kMaxRing];
    ScanState states[kMaxMachines];
    ScanMachine machines[kMaxMachines];
};

// The scratch itself.  In HOST code this name is a host-side stand-in, not a device address: host code below only ever
// hands the address cudaGetSymbolAddress returns for the current device to a kernel.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(Scratch)];

struct MachineTable {  // passed to the set-up kernel by value (kernel parameter space, < 4 KB)
    ScanMachine m[kMaxMachines];
};

// Re-initialise the scratch for M machines: machine table, rings (RollingAverage.init pre-fill of the long-term ring,
// reference src/structures/RollingAverage.zig:19-25: written_count = len, last_avg = sum_i init*(1/len) accumulated
// sequentially in f64; everything else zero) and the scan state.
__global__ void noalloc_setup_kernel(Scratch *sc, const MachineTable tab, int M, int pitch) {
    const int m = blockIdx.x;
    if (m >= M) return;
    const ScanMachine mc = tab.m[m];
    double *ring = sc->rings + (size_t)m * pitch;
    for (int i = threadIdx.x; i < pitch; i += blockDim.x) ring[i] = (mc.has_init && i < mc.n_long) ? mc.init_avg : 0.0;
    if (threadIdx.x == 0) {
        sc->machines[m] = mc;
        ScanState st;
        memset(&st, 0, sizeof(st));
        if (mc.has_init) {
            st.long_cnt = mc.n_long;
            const double scalar = 1.0 / (double)mc.n_long;
            double acc = 0.0, sum = 0.0;
            for (int i = 0; i < mc.n_long; ++i) { acc += mc.init_avg * scalar; sum += mc.init_avg; }
            st.last_avg = acc;
            st.long_sum = sum;
            st.has_last = 1;
        }
        sc->states[m] = st;
    }
}

__global__ void noalloc_counts_kernel(const Scratch *sc, unsigned long long *n_segments, int M) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m < M) n_segments[m] = sc->states[m].n_segments;
}

// device address of the scratch on the CURRENT device, looked up once per device
Scratch *scratch_on_current_device() {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static Scratch *ptr[kMaxDevices] = {};
    int dev = 0;
    FVAD_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) throw Error(FVAD_ERR_INVALID_ARGUMENT, "device ordinal beyond the no-alloc table");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            throw Error(FVAD_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
        ptr[dev] = static_cast<Scratch *>(p);
    }
    return ptr[dev];
}

void scan_noalloc(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate, size_t fft_size,
                  const float *d_band_min, const float *d_ratio, size_t n_frames, unsigned char *d_met, fvad_segment *d_segments,
                  size_t max_segments, unsigned long long *d_n_segments, cudaStream_t stream) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        throw Error(FVAD_ERR_NO_DEVICE, "no CUDA device: libfvad_cuda has no CPU path");
    }
    if (!cfgs || !n_machines || !d_n_segments || (n_frames && (!d_band_min || !d_ratio)) || (max_segments && !d_segments))
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
    if (fft_size == 0 || sample_rate == 0) throw Error(FVAD_ERR_INVALID_ARGUMENT, "sample_rate and fft_size must be positive");
    if (n_machines > (size_t)kMaxMachines)
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "fvad_vad_machine_scan_noalloc: at most " + std::to_string(kMaxMachines) +
                                                   " machines per call (FVAD_NOALLOC_MAX_MACHINES); use fvad_vad_machine_scan");
    if (n_frames > (size_t)INT_MAX || max_segments > (size_t)INT_MAX || fft_size > (size_t)INT_MAX)
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "fvad_vad_machine_scan_noalloc: n_frames / max_segments / fft_size beyond 2^31-1");
    const int M = (int)n_machines;
    MachineTable tab;
    memset(&tab, 0, sizeof(tab));
    int pitch = 0, fast = 1;
    for (int m = 0; m < M; ++m) {
        ScanMachine d = derive_machine(cfgs[m], sample_rate, fft_size);
        d.band = m;  // machine m scans row m of band_min
        const long long need = (long long)d.n_long + d.n_short + d.n_ratio;
        if (need > kMaxRing)
            throw Error(FVAD_ERR_INVALID_ARGUMENT,
                        "fvad_vad_machine_scan_noalloc: machine " + std::to_string(m) + " needs " + std::to_string(need) +
                            " ring slots (long + short + ratio windows in frames), the static scratch holds " +
                            std::to_string(kMaxRing) + " (FVAD_NOALLOC_MAX_RING); use fvad_vad_machine_scan");
        pitch = pitch > (int)need ? pitch : (int)need;
        if (d.n_long < 2 * 256 + 2 || d.n_short + d.n_ratio > 64) fast = 0;  // vadscan.cu fast-path limits
        tab.m[m] = d;
    }
    Scratch *sc = scratch_on_current_device();

    noalloc_setup_kernel<<<(unsigned)M, 256, 0, stream>>>(sc, tab, M, pitch);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
    if (n_frames) {
        ScanArgs ca;
        ca.band = d_band_min; ca.ratio = d_ratio; ca.machines = sc->machines; ca.states = sc->states;
        ca.rings = sc->rings; ca.ring_pitch = pitch; ca.segments = d_segments; ca.max_segments = (int)max_segments;
        ca.met = d_met; ca.met_per_machine = 1; ca.S = 1; ca.M = M;
        ca.n_frames = (int)n_frames; ca.max_frames = (int)n_frames;
        ca.first_frame = 0; ca.fft_size = (int)fft_size; ca.fast = fast;
        launch_vadscan(ca, stream);
    }
    noalloc_counts_kernel<<<1, 32, 0, stream>>>(sc, d_n_segments, M);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
    FVAD_CUDA_CHECK(cudaStreamSynchronize(stream));
}

}  // namespace
}  // namespace fvad

extern "C" FVAD_API int fvad_vad_machine_scan_noalloc(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate,
                                                      size_t fft_size, const float *d_band_min, const float *d_ratio, size_t n_frames,
                                                      uint8_t *d_met, fvad_segment *d_segments, size_t max_segments,
                                                      uint64_t *d_n_segments, void *cuda_stream) {
    using namespace fvad;
    try {
        scan_noalloc(cfgs, n_machines, sample_rate, fft_size, d_band_min, d_ratio, n_frames, d_met, d_segments, max_segments,
                     reinterpret_cast<unsigned long long *>(d_n_segments), static_cast<cudaStream_t>(cuda_stream));
        return FVAD_OK;
    } catch (const Error &e) {
        set_last_error(e.what());
        return e.code;
    } catch (const std::exception &e) {
        set_last_error(e.what());
        return FVAD_ERR_INVALID_ARGUMENT;
    }
}
