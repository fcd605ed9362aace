// nn.cuh -- NSNet2 network backends.  Replaces `onnx_instance.run()` (reference src/NSNet2.zig:220,
// lib/onnxruntime.zig/src/onnxruntime.zig:200-213): FC 161->400, GRU 400 x2, FC 400->600 ReLU,
// FC 600->600 ReLU, FC 600->161 sigmoid on a [54,161] window per chunk, zero initial GRU state.
#pragma once
#include <memory>
#include <vector>

#include "common.cuh"

namespace fvad {

// Host copy of the weights in the ONNX layouts of fvad_nsnet2_weights.
struct HostWeights {
    int n_bins = 161, hidden = 400, fc2 = 600, fc3 = 600, fc1_relu = 0;
    std::vector<float> fc1_w, fc1_b, gru_W[2], gru_R[2], gru_Wb[2], gru_Rb[2], fc2_w, fc2_b, fc3_w, fc3_b, fc4_w, fc4_b;
    void from_struct(const fvad_nsnet2_weights &w);
    void from_file(const char *path);  // FVADNS2 blob
    void validate() const;
};

// One span of work: every chunk j of every stream-channel sc is one 54-step sequence whose input
// rows are feat[sc][50j .. 50j+53] (overlapping windows: rows 50j..50j+3 are the 4 context frames)
// and whose output rows 4..53 land in gains[sc][50j .. 50j+49].
struct NnSpan {
    const float *feat;  // [SC][4+50K][161], or nullptr when the backend wants_feat16() and nothing else needs the f32 rows
    float *gains;       // [SC][50K][161]
    int SC, K;
    const void *feat16; // the same rows as fp16 [SC][4+50K][176], or nullptr (only a backend with wants_feat16() gets them)
    cudaEvent_t after_rnn = nullptr;  // recorded on the stream once the span's last recurrent launch is queued (or at the
                                      // end of forward): the engine starts the previous span's VAD scan there
};

class NnBackend {
public:
    virtual ~NnBackend() = default;
    virtual void forward(const NnSpan &span, cudaStream_t s) = 0;
    virtual const char *name() const = 0;
    virtual bool wants_feat16() const { return false; }  // reads NnSpan::feat16 instead of NnSpan::feat
    // CUDA-event ms and launch count per layer kernel {pack, fc1, W1, gru1, W2, gru2, fc2, fc3, fc4} since the last call
    virtual void take_layer_ms(double *ms9, unsigned *launches9) { for (int i = 0; i < 9; ++i) { ms9[i] = 0; launches9[i] = 0; } }
};

std::unique_ptr<NnBackend> make_nn_f32(const HostWeights &w);
std::unique_ptr<NnBackend> make_nn_tc(const HostWeights &w, bool split3);  // nn_tc.cu

}  // namespace fvad
