// eval.cu -- the Evaluator on the device: scores every (stream, VAD-machine) segment list against the
// stream's labelled reference segments in one launch (BASELINE config 5: 256 machine configs x 20 streams
// = 5 120 lists, which never leave HBM).
//
// Replaces (reference):
//   Evaluator.initAndRun              (src/Evaluator.zig:90-156)         sort by start, all-pairs overlap matching
//   SpeechSegment.overlapWith/duration (src/Evaluator/SpeechSegment.zig:18-27)
//   statistics.fromEvaluator          (src/Evaluator/statistics.zig:85-114)
//   calcFalsePositiveSec / calcTruePositiveSec / calcFalseNegativeSec / extrudeSegments / calcOverlapMany
//                                     (statistics.zig:191-283)
//   SimulationInstance.storeResult    (src/simulator/SimulationInstance.zig:227-255) sample index -> seconds as
//                                     f32(sample) / f32(sample_rate)
//
// One thread per list.  All sums are f32 and run in the reference's order (VAD segments by start, reference
// segments by start), so the statistics are the reference's to the bit given the same segments.  The
// reference clones and extrudes the matched reference segments of each VAD segment; here that is a single
// streaming pass with one pending segment (the gap fill only ever looks at a segment and its successor, and
// only the first/last matched segments are extruded), so no per-thread storage is needed.
#include "engine.cuh"

namespace fvad {

namespace {

struct EvalSeg { float from, to; };

__device__ __forceinline__ float seg_overlap(EvalSeg a, EvalSeg b) {  // SpeechSegment.zig:22-27 (may be negative)
    return fminf(a.to, b.to) - fmaxf(a.from, b.from);
}

// VAD side accessor: either seconds given directly or sample indices of a machine's segment list
struct VadList {
    const fvad_segment *samples;   // device, or nullptr
    const fvad_eval_segment *secs; // device, or nullptr
    unsigned n;
    float sample_rate;
    __device__ __forceinline__ EvalSeg get(unsigned i) const {
        if (samples) return EvalSeg{__fdiv_rn((float)samples[i].sample_from, sample_rate), __fdiv_rn((float)samples[i].sample_to, sample_rate)};
        return EvalSeg{secs[i].from_sec, secs[i].to_sec};
    }
};

__device__ void evaluate_list(const VadList &vad, const fvad_eval_segment *__restrict__ ref, unsigned n_ref,
                              const fvad_eval_config &cfg, fvad_eval_stats &st) {
    float tp_sum = 0.0f, fp_sum = 0.0f, fn_sum = 0.0f, p_sum = 0.0f;
    // ---- VAD segments: false / true positive seconds (statistics.zig:88-94) ----
    for (unsigned i = 0; i < vad.n; ++i) {
        const EvalSeg v = vad.get(i);
        float overlap = 0.0f;
        bool have = false;
        EvalSeg pend{0.0f, 0.0f};
        for (unsigned j = 0; j < n_ref; ++j) {
            const EvalSeg r{ref[j].from_sec, ref[j].to_sec};
            if (!(seg_overlap(v, r) > 0.0f)) continue;       // findOverlapping, SpeechSegment.zig:44-56
            EvalSeg cur = r;
            if (!have) {
                cur.from -= cfg.extrude_start;                  // extrudeSegments: first.from_sec -= extrude_start
            } else {
                if (cur.from - pend.to <= cfg.fill_gaps) pend.to = cur.from;  // bridge the gap to the successor
                overlap += fmaxf(0.0f, seg_overlap(v, pend));  // calcOverlapMany
            }
            pend = cur;
            have = true;
        }
        if (have) {
            pend.to += cfg.extrude_end;                         // last.to_sec += extrude_end
            overlap += fmaxf(0.0f, seg_overlap(v, pend));
        }
        const float dur = v.to - v.from;
        const float fp = fmaxf(0.0f, dur - overlap);            // calcFalsePositiveSec
        const float tp = fmaxf(0.0f, dur - fp);                 // calcTruePositiveSec
        fp_sum += fp;
        tp_sum += tp;
        p_sum += tp;
    }
    // ---- reference segments: false negative seconds (statistics.zig:96-102) ----
    for (unsigned j = 0; j < n_ref; ++j) {
        const EvalSeg r{ref[j].from_sec, ref[j].to_sec};
        const float dur = r.to - r.from;
        if (dur < cfg.ignore_shorter_than_sec) continue;
        float ov = 0.0f;
        for (unsigned i = 0; i < vad.n; ++i) {
            const float o = seg_overlap(r, vad.get(i));
            if (o > 0.0f) ov += o;                              // calcOverlapWithMatches over the matched segments
        }
        const float fn = fmaxf(0.0f, dur - ov);
        fn_sum += fn;
        p_sum += fn;
    }
    st.total_positives_sec = p_sum;
    st.true_positives_sec = tp_sum;
    st.false_positives_sec = fp_sum;
    st.false_negatives_sec = fn_sum;
    st.true_positive_rate = __fdiv_rn(tp_sum, p_sum);
    st.false_negative_rate = __fdiv_rn(fn_sum, p_sum);
    st.false_discovery_rate = __fdiv_rn(fp_sum, fp_sum + tp_sum);
    st.precision = __fdiv_rn(tp_sum, tp_sum + fp_sum);
    const float b2 = __fmul_rn(cfg.f_score_beta, cfg.f_score_beta);  // statistics.zig:175-177
    st.f_score = __fdiv_rn(__fmul_rn(1.0f + b2, __fmul_rn(st.precision, st.true_positive_rate)),
                           __fadd_rn(__fmul_rn(b2, st.precision), st.true_positive_rate));
    st.fm_index = sqrtf(__fmul_rn(st.precision, st.true_positive_rate));
}

__global__ void eval_batch_kernel(const fvad_segment *segments, const ScanState *states, int max_segments, int S, int M,
                                  float sample_rate, const fvad_eval_segment *refs, const unsigned long long *ref_off,
                                  fvad_eval_config cfg, fvad_eval_stats *out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;  // (stream, machine)
    if (t >= S * M) return;
    const int s = t / M;
    VadList vl{segments + (size_t)t * max_segments, nullptr, min(states[t].n_segments, (unsigned)max_segments), sample_rate};
    evaluate_list(vl, refs + ref_off[s], (unsigned)(ref_off[s + 1] - ref_off[s]), cfg, out[t]);
}

__global__ void eval_single_kernel(const fvad_eval_segment *vad, unsigned n_vad, const fvad_eval_segment *ref, unsigned n_ref,
                                   fvad_eval_config cfg, fvad_eval_stats *out) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        VadList vl{nullptr, vad, n_vad, 1.0f};
        evaluate_list(vl, ref, n_ref, cfg, *out);
    }
}

// std.mem.sort(..., sortByStart) is a stable sort on from_sec (Evaluator.zig:110-111)
void sort_by_start(std::vector<fvad_eval_segment> &v, size_t lo, size_t hi) {
    std::stable_sort(v.begin() + (long)lo, v.begin() + (long)hi,
                     [](const fvad_eval_segment &a, const fvad_eval_segment &b) { return a.from_sec < b.from_sec; });
}

}  // namespace

void Batch::evaluate(const fvad_eval_segment *refs, const size_t *ref_offsets, const fvad_eval_config &cfg, fvad_eval_stats *out) {
    if (!ref_offsets || !out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
    FVAD_CUDA_CHECK(cudaSetDevice(opt_.device));
    const size_t n_total = ref_offsets[S_];
    if (n_total && !refs) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL reference segments");
    std::vector<fvad_eval_segment> r(refs, refs + n_total);
    std::vector<unsigned long long> off((size_t)S_ + 1);
    for (int s = 0; s <= S_; ++s) {
        if (s && ref_offsets[s] < ref_offsets[s - 1]) throw Error(FVAD_ERR_INVALID_ARGUMENT, "reference offsets must be non-decreasing");
        off[(size_t)s] = ref_offsets[s];
    }
    for (int s = 0; s < S_; ++s) sort_by_start(r, ref_offsets[s], ref_offsets[s + 1]);
    DevBuf<fvad_eval_segment> d_ref;
    DevBuf<unsigned long long> d_off;
    DevBuf<fvad_eval_stats> d_out;
    const size_t SM = (size_t)S_ * M_;
    d_ref.ensure(std::max<size_t>(n_total, 1));
    d_off.ensure(off.size());
    d_out.ensure(SM);
    cudaStream_t s = stream_;
    if (n_total) FVAD_CUDA_CHECK(cudaMemcpyAsync(d_ref.p, r.data(), n_total * sizeof(fvad_eval_segment), cudaMemcpyHostToDevice, s));
    FVAD_CUDA_CHECK(cudaMemcpyAsync(d_off.p, off.data(), off.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice, s));
    FVAD_CUDA_CHECK(cudaStreamWaitEvent(s, scan_done_, 0));
    eval_batch_kernel<<<(unsigned)((SM + 127) / 128), 128, 0, s>>>(d_segments_.p, d_states_.p, max_segments_, S_, M_, (float)sample_rate_,
                                                                  d_ref.p, d_off.p, cfg, d_out.p);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
    FVAD_CUDA_CHECK(cudaMemcpyAsync(out, d_out.p, SM * sizeof(fvad_eval_stats), cudaMemcpyDeviceToHost, s));
    FVAD_CUDA_CHECK(cudaStreamSynchronize(s));
}

void evaluate_lists(const fvad_eval_segment *vad, size_t n_vad, const fvad_eval_segment *ref, size_t n_ref,
                    const fvad_eval_config &cfg, fvad_eval_stats *out) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw Error(FVAD_ERR_NO_DEVICE, "no CUDA device: libfvad_cuda has no CPU path");
    if ((n_vad && !vad) || (n_ref && !ref) || !out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
    std::vector<fvad_eval_segment> v(vad, vad + n_vad), r(ref, ref + n_ref);
    sort_by_start(v, 0, n_vad);
    sort_by_start(r, 0, n_ref);
    DevBuf<fvad_eval_segment> dv, dr;
    DevBuf<fvad_eval_stats> d_out;
    dv.ensure(std::max<size_t>(n_vad, 1));
    dr.ensure(std::max<size_t>(n_ref, 1));
    d_out.ensure(1);
    if (n_vad) FVAD_CUDA_CHECK(cudaMemcpy(dv.p, v.data(), n_vad * sizeof(fvad_eval_segment), cudaMemcpyHostToDevice));
    if (n_ref) FVAD_CUDA_CHECK(cudaMemcpy(dr.p, r.data(), n_ref * sizeof(fvad_eval_segment), cudaMemcpyHostToDevice));
    eval_single_kernel<<<1, 32>>>(dv.p, (unsigned)n_vad, dr.p, (unsigned)n_ref, cfg, d_out.p);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
    FVAD_CUDA_CHECK(cudaMemcpy(out, d_out.p, sizeof(fvad_eval_stats), cudaMemcpyDeviceToHost));
}

}  // namespace fvad
