// weights.cpp -- NSNet2 weight container: validation, in-memory struct import, FVADNS2 blob reader.
// Stands where the reference opens its ONNX session (src/NSNet2.zig:53-66,
// lib/onnxruntime.zig/src/onnxruntime.zig:76-142); model-load failures map to FVAD_ERR_MODEL
// (the reference's error.OnnxError).
#include <cstdio>
#include <cstring>

#include "nn.cuh"

namespace fvad {

static thread_local std::string t_last_error;
void set_last_error(const std::string &msg) { t_last_error = msg; }
const std::string &last_error() { return t_last_error; }

static void take(std::vector<float> &dst, const float *src, size_t n, const char *name) {
    if (!src) throw Error(FVAD_ERR_MODEL, std::string("weights: missing tensor ") + name);
    dst.assign(src, src + n);
}

void HostWeights::from_struct(const fvad_nsnet2_weights &w) {
    n_bins = w.n_bins; hidden = w.hidden; fc2 = w.fc2; fc3 = w.fc3; fc1_relu = w.fc1_relu;
    validate();
    const size_t H = (size_t)hidden, B = (size_t)n_bins, F2 = (size_t)fc2, F3 = (size_t)fc3;
    take(fc1_w, w.fc1_w, H * B, "fc1_w"); take(fc1_b, w.fc1_b, H, "fc1_b");
    take(gru_W[0], w.gru1_W, 3 * H * H, "gru1_W"); take(gru_R[0], w.gru1_R, 3 * H * H, "gru1_R");
    take(gru_Wb[0], w.gru1_Wb, 3 * H, "gru1_Wb"); take(gru_Rb[0], w.gru1_Rb, 3 * H, "gru1_Rb");
    take(gru_W[1], w.gru2_W, 3 * H * H, "gru2_W"); take(gru_R[1], w.gru2_R, 3 * H * H, "gru2_R");
    take(gru_Wb[1], w.gru2_Wb, 3 * H, "gru2_Wb"); take(gru_Rb[1], w.gru2_Rb, 3 * H, "gru2_Rb");
    take(fc2_w, w.fc2_w, F2 * H, "fc2_w"); take(fc2_b, w.fc2_b, F2, "fc2_b");
    take(fc3_w, w.fc3_w, F3 * F2, "fc3_w"); take(fc3_b, w.fc3_b, F3, "fc3_b");
    take(fc4_w, w.fc4_w, B * F3, "fc4_w"); take(fc4_b, w.fc4_b, B, "fc4_b");
}

void HostWeights::validate() const {
    // the kernels are specialised for the NSNet2-baseline shapes the reference ships
    if (n_bins != kBins || hidden != kHidden || fc2 != 600 || fc3 != 600)
        throw Error(FVAD_ERR_MODEL, "weights: expected NSNet2 baseline shapes 161/400/600/600");
}

static void read_exact(FILE *f, void *dst, size_t bytes, const char *path) {
    if (fread(dst, 1, bytes, f) != bytes) {
        fclose(f);
        throw Error(FVAD_ERR_MODEL, std::string("weights: truncated file ") + path);
    }
}

void HostWeights::from_file(const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) throw Error(FVAD_ERR_MODEL, std::string("weights: cannot open ") + path);
    char magic[8];
    read_exact(f, magic, 8, path);
    if (memcmp(magic, "FVADNS2\0", 8) != 0) {
        fclose(f);
        throw Error(FVAD_ERR_MODEL, std::string("weights: ") + path +
                                        " is not an FVADNS2 blob (convert .onnx with tools/onnx_to_blob.py)");
    }
    uint32_t hdr[6];
    read_exact(f, hdr, sizeof(hdr), path);
    if (hdr[0] != 1) { fclose(f); throw Error(FVAD_ERR_MODEL, "weights: unsupported blob version"); }
    fc1_relu = (int)hdr[1]; n_bins = (int)hdr[2]; hidden = (int)hdr[3]; fc2 = (int)hdr[4]; fc3 = (int)hdr[5];
    try { validate(); } catch (...) { fclose(f); throw; }
    const size_t H = (size_t)hidden, B = (size_t)n_bins, F2 = (size_t)fc2, F3 = (size_t)fc3;
    auto rd = [&](std::vector<float> &v, size_t n) { v.resize(n); read_exact(f, v.data(), n * sizeof(float), path); };
    rd(fc1_w, H * B); rd(fc1_b, H);
    for (int l = 0; l < 2; ++l) { rd(gru_W[l], 3 * H * H); rd(gru_R[l], 3 * H * H); rd(gru_Wb[l], 3 * H); rd(gru_Rb[l], 3 * H); }
    rd(fc2_w, F2 * H); rd(fc2_b, F2); rd(fc3_w, F3 * F2); rd(fc3_b, F3); rd(fc4_w, B * F3); rd(fc4_b, B);
    fclose(f);
}

}  // namespace fvad
