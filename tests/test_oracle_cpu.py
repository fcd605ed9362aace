"""CPU tests (-m "not gpu"): the oracle against the reference's own kissfft, the reference's unit-test
known answers, independent re-derivations (numpy / torch fp64), and the committed golden
fixtures.  Nothing here needs a GPU."""
import ctypes as C
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


# ---------------------------------------------------------------- FFT: restatement == reference kissfft
@pytest.mark.parametrize("n", [320, 1024, 64, 160, 512, 30])
def test_oracle_fft_is_bit_exact_to_reference_kissfft(po, n):
    if not po.use_reference_kissfft(True):
        pytest.skip("oracle/_ref/libkissfft_ref.so not built (reference checkout absent)")
    try:
        rng = np.random.default_rng(n)
        x = rng.standard_normal((5, n)).astype(np.float32)
        w = rng.uniform(0.1, 1, n).astype(np.float32)
        ref_f = po.rfft(x, w)
        ref_i = po.irfft(ref_f, n)
    finally:
        po.use_reference_kissfft(False)
    got_f = po.rfft(x, w)
    got_i = po.irfft(ref_f, n)
    assert np.array_equal(got_f.view(np.uint32), ref_f.view(np.uint32)), "forward rFFT differs from kissfft bitwise"
    assert np.array_equal(got_i.view(np.uint32), ref_i.view(np.uint32)), "inverse rFFT differs from kissfft bitwise"


@pytest.mark.parametrize("n", [320, 1024])
def test_oracle_fft_against_numpy(po, n):
    rng = np.random.default_rng(7)
    x = rng.standard_normal((4, n)).astype(np.float32)
    X = po.rfft(x)
    ref = np.fft.rfft(x.astype(np.float64), axis=-1)
    assert np.abs(X - ref).max() / np.abs(ref).max() < 2e-6  # kissfft's own test wants 90 dB SNR (testkiss.py)
    back = po.irfft(X, n)  # unnormalised inverse: n * x
    assert np.abs(back / n - x).max() < 5e-6


def test_windows_follow_window_fn_zig(po):
    n = 1024
    w = po.hann_periodic(n)
    ref = 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(n) / n)  # window_fn.zig:22-28,51-68
    assert np.abs(w - ref).max() < 2e-7
    ws = po.nsnet2_window()  # sqrt(symmetric Hann(320)), NSNet2.zig:384-396
    ref = np.sqrt(0.5 - 0.5 * np.cos(2 * np.pi * np.arange(320) / 319))
    assert np.abs(ws - ref).max() < 1e-6  # 0.5-0.5cos computed in f32 (window_fn.zig:30-41), then sqrt: cancellation near the ends
    assert abs(ws[0]) < 1e-7 and abs(ws[-1]) < 5e-4
    norm = po.lib().orc_window_norm_factor(po.f32p(w), n)  # N / sum(w), window_fn.zig:8-16
    assert abs(norm - 2.0) < 1e-5


def test_resample_and_rms(po):
    L = po.lib()
    x = np.arange(30, dtype=np.float32)
    out = np.zeros(10, np.float32)
    L.orc_downsample(po.f32p(x), 30, po.f32p(out), 3)  # resample.zig:9-29: every 3rd sample, no filter
    assert np.array_equal(out, x[0::3])
    y = np.array([3.0, 6.0, 0.0], np.float32)
    up = np.zeros(9, np.float32)
    last = L.orc_upsample(po.f32p(y), 3, po.f32p(up), C.c_float(0.0), 3)  # resample.zig:32-79
    assert np.allclose(up, [1, 2, 3, 4, 5, 6, 4, 2, 0], atol=1e-6)  # true sample at 3m+2, lerp before it
    assert last == 0.0
    v = np.full(24000, 0.5, np.float32)
    assert abs(L.orc_rms_volume(po.f32p(v), v.size) - 0.5) < 1e-4  # audio_utils.zig:14-24 (sequential f32 sum)
    assert L.orc_freq_to_bin(500.0, 48000, 1024) == 11 and L.orc_freq_to_bin(2000.0, 48000, 1024) == 43
    assert L.orc_freq_to_bin(25000.0, 48000, 1024) < 0 and L.orc_freq_to_bin(-1.0, 48000, 1024) < 0


# ---------------------------------------------------------------- reference unit-test known answers
def _meta_run(po, pushes):
    L = po.lib()
    m = po.Meta()
    for kw, weight in pushes:
        r = po.MetaResult()
        for k, v in kw.items():
            setattr(r, "has_" + {"ratio": "ratio", "vmin": "min", "vmax": "max"}[k], 1)
            setattr(r, k, v)
        L.orc_meta_push(C.byref(m), C.byref(r), C.c_float(weight))
    return L.orc_meta_to_result(C.byref(m))


def test_vadmetadata_known_answers(po):
    """src/AudioPipeline/VADMetadata.zig:70-110, verbatim."""
    assert abs(_meta_run(po, [({"vmin": 100}, 1), ({"vmin": 80}, 1), ({"vmin": 90}, 1)]).vmin - 80.0) < 1e-3
    assert abs(_meta_run(po, [({"vmax": 80}, 1), ({"vmax": 100}, 1), ({"vmax": 90}, 1)]).vmax - 100.0) < 1e-3
    assert abs(_meta_run(po, [({"ratio": 0.9}, 1), ({"ratio": 0.8}, 1)]).ratio - 0.85) < 1e-3
    assert abs(_meta_run(po, [({"ratio": 1.0}, 1), ({"ratio": 0.0}, 9)]).ratio - 0.1) < 1e-3
    assert _meta_run(po, []).has_ratio == 0


def test_statistics_false_positive_known_answers(po):
    """src/Evaluator/statistics.zig:286-360, verbatim: FP = 0 s and 3 s."""
    L = po.lib()
    refs = (po.EvalSeg * 2)(po.EvalSeg(2, 3), po.EvalSeg(4, 5))
    cfg = po.EvalCfg(2, 2, 2, 0, 0.7)
    v1, v2 = po.EvalSeg(1, 6), po.EvalSeg(1, 10)
    assert abs(L.orc_eval_false_positive_sec(C.byref(v1), refs, 2, C.byref(cfg)) - 0.0) < 1e-3
    assert abs(L.orc_eval_false_positive_sec(C.byref(v2), refs, 2, C.byref(cfg)) - 3.0) < 1e-3


def test_evaluator_rates(po):
    """Evaluator.zig:90-156 + statistics.zig:85-114 on a hand-checked case."""
    L = po.lib()
    vad = (po.EvalSeg * 2)(po.EvalSeg(10, 20), po.EvalSeg(50, 55))
    ref = (po.EvalSeg * 2)(po.EvalSeg(12, 18), po.EvalSeg(80, 90))
    cfg = po.EvalCfg(0, 0, 0, 0, 0.7)
    st = po.EvalStats()
    L.orc_evaluate(vad, 2, ref, 2, C.byref(cfg), C.byref(st))
    # vad#1 overlaps ref#1 for 6 s (4 s FP); vad#2 is 5 s of FP; ref#2 is 10 s of FN
    assert abs(st.tp_sec - 6) < 1e-3 and abs(st.fp_sec - 9) < 1e-3 and abs(st.fn_sec - 10) < 1e-3
    assert abs(st.p_sec - 16) < 1e-3
    assert abs(st.tpr - 6 / 16) < 1e-4 and abs(st.ppv - 6 / 15) < 1e-4


# ---------------------------------------------------------------- RollingAverage / VADMachine
def test_rolling_average(po):
    L = po.lib()
    r = L.orc_rolling_new(3, 0, 0.0)
    assert L.orc_rolling_push(r, 3.0) == pytest.approx(3.0)       # divides by written_count, not len
    assert L.orc_rolling_push(r, 6.0) == pytest.approx(4.5)
    assert L.orc_rolling_push(r, 0.0) == pytest.approx(3.0)
    assert L.orc_rolling_push(r, 9.0) == pytest.approx(5.0)       # 9,6,0 in array order
    L.orc_rolling_free(r)
    r = L.orc_rolling_new(4, 1, 0.005)                             # pre-filled: written = len
    assert L.orc_rolling_push(r, 0.0) == pytest.approx(0.00375)
    L.orc_rolling_free(r)


def _run_machine(po, bands, ratios=None, sr=48000, **cfg):
    L = po.lib()
    c = po.VadMachineConfig.default(**cfg)
    m = L.orc_vadm_new(C.byref(c), sr, 1, 1024)
    mets, states = [], []
    for i, b in enumerate(bands):
        v = np.array([b], np.float32)
        sn = C.c_uint64()
        L.orc_vadm_run(m, po.f32p(v), i * 1024, 1, C.c_float(1.0 if ratios is None else ratios[i]), C.byref(sn))
        mets.append(L.orc_vadm_last_met(m))
        states.append(L.orc_vadm_state(m))
    p = C.POINTER(po.Segment)()
    n = L.orc_vadm_segments(m, C.byref(p))
    segs = [(p[i].sample_from, p[i].sample_to, p[i].avg_channel_vol_ratio, p[i].vad_met_sec) for i in range(n)]
    L.orc_vadm_free(m)
    return mets, states, segs


def test_vad_machine_transitions(po):
    """VADMachine.zig:138-325 on a hand-built band-energy track."""
    quiet, loud = 0.004, 1.0
    bands = [quiet] * 50 + [loud] * 60 + [quiet] * 120
    mets, states, segs = _run_machine(po, bands)
    # short-term mean (9 frames) crosses 10 x 0.005 on the first loud frame
    first = mets.index(1)
    assert first == 50
    assert states[first] == 1                       # closed -> opening
    assert states[first + 9] == 1 and states[first + 10] == 2   # open once idx - start >= 9600 (10 frames)
    last_met = max(i for i, m in enumerate(mets) if m)
    assert states[last_met + 1] == 3                # first !met frame: closing, end = that frame
    end_frame = last_met + 1
    assert states[end_frame + 93] == 3 and states[end_frame + 94] == 0  # gap >= 96000 samples -> 94 frames
    assert len(segs) == 1
    start, end = 50 * 1024, end_frame * 1024
    assert segs[0][0] == start - min(96000, start) and segs[0][1] == end + 96000
    assert segs[0][2] == pytest.approx(1.0)
    # a burst shorter than min_vad_duration (0.7 s ~ 33 frames) is aborted; an open segment at the end is not flushed
    mets, states, segs = _run_machine(po, [quiet] * 50 + [loud] * 14 + [quiet] * 130 + [loud] * 60)
    assert segs == [] and states[-1] == 2
    # channel ratio condition (ratio mean over 23 frames must exceed 0.5)
    mets, _, segs = _run_machine(po, bands, ratios=[0.2] * len(bands))
    assert sum(mets) == 0 and segs == []


# ---------------------------------------------------------------- NSNet2 network vs torch fp64
def test_oracle_nn_matches_torch(po, wd):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(3)
    feat = rng.uniform(-8, 2, (54, 161)).astype(np.float32)
    got = po.nn_forward(wd, feat)
    got64 = po.nn_forward(wd, feat, f64=True)
    H = 400
    t = lambda a: torch.tensor(np.asarray(a, np.float64))
    x = t(feat) @ t(wd["fc1_w"]).T + t(wd["fc1_b"])
    for l in (1, 2):
        gru = torch.nn.GRU(H, H, batch_first=False).double()
        W, R, Wb, Rb = (wd[f"gru{l}_{k}"] for k in ("W", "R", "Wb", "Rb"))
        perm = np.r_[H:2 * H, 0:H, 2 * H:3 * H]  # ONNX z,r,h -> torch r,z,n
        with torch.no_grad():
            gru.weight_ih_l0.copy_(t(W[perm])); gru.weight_hh_l0.copy_(t(R[perm]))
            gru.bias_ih_l0.copy_(t(Wb[perm])); gru.bias_hh_l0.copy_(t(Rb[perm]))
            x, _ = gru(x[:, None, :])
        x = x[:, 0, :]
    with torch.no_grad():
        x = torch.relu(x @ t(wd["fc2_w"]).T + t(wd["fc2_b"]))
        x = torch.relu(x @ t(wd["fc3_w"]).T + t(wd["fc3_b"]))
        ref = torch.sigmoid(x @ t(wd["fc4_w"]).T + t(wd["fc4_b"])).numpy()
    assert np.abs(got64 - ref).max() < 1e-12
    assert np.abs(got - ref).max() < 2e-6


# ---------------------------------------------------------------- pipeline semantics
def test_pipeline_chunking_and_errors(po, wd, fv):
    x, _ = fv.synth.onboard_stream(3, 3.3, 16000, 1)
    x48 = fv.synth.expand_16k_to_48k(x)
    with pytest.raises(ValueError, match="InvalidSampleRate"):
        po.Pipeline(wd, 44100, 1, allow_any_rate=False)     # VADPipeline.zig:55-58
    with pytest.raises(ValueError, match="InvalidSampleRate"):
        po.Pipeline(wd, 16000, 1, allow_any_rate=False)     # the reference only takes 48 kHz
    a = po.Pipeline(wd, 48000, 1, keep_denoised=True)
    assert a.push_samples(x48) == 0
    band, ratio, met = a.frames()
    n_chunks = x48.shape[1] // 24000                        # trailing partial chunk is never processed (VADPipeline.zig:150)
    assert len(band) == n_chunks * 24000 // 1024 and a.denoised(0).size == n_chunks * 24000
    # ragged pushes give the identical result (ring + chunk carry)
    b = po.Pipeline(wd, 48000, 1, keep_denoised=True)
    cuts = [0, 777, 24000, 24001, 70000, x48.shape[1]]
    for i, j in zip(cuts[:-1], cuts[1:]):
        assert b.push_samples(x48[:, i:j]) == i
    assert np.array_equal(a.denoised(0), b.denoised(0))
    assert np.array_equal(a.frames()[0], b.frames()[0])
    # 16 kHz native (documented deviation) sees the same decimated signal: identical gains
    c = po.Pipeline(wd, 16000, 1, keep_nn=True)
    c.push_samples(x)
    d = po.Pipeline(wd, 48000, 1, keep_nn=True)
    d.push_samples(x48)
    assert np.array_equal(c.nn_taps()[1], d.nn_taps()[1])
    # mono: volume ratio is identically 1
    assert np.all(ratio == 1.0)


def test_first_chunk_context_rows_are_zero(po, wd, fv):
    """NSNet2.zig:79,188-203: the 4 context feature rows of the first chunk are 0.0, not -12."""
    L = po.lib()
    ws, keep = po.make_weights_struct(wd)
    h = L.orc_nsnet2_new(16000, C.byref(ws))
    x, _ = fv.synth.onboard_stream(5, 1.0, 16000, 1)
    out = np.empty(8000, np.float32)
    assert L.orc_nsnet2_denoise(h, po.f32p(np.ascontiguousarray(x[0, :8000])), 8000, po.f32p(out)) == 0
    f = np.ctypeslib.as_array(L.orc_nsnet2_last_features(h), (54, 161)).copy()
    assert np.all(f[:4] == 0.0) and np.all(f[4:] >= -12.0)
    first_rows = f[50:54].copy()
    assert L.orc_nsnet2_denoise(h, po.f32p(np.ascontiguousarray(x[0, 8000:16000])), 8000, po.f32p(out)) == 0
    f2 = np.ctypeslib.as_array(L.orc_nsnet2_last_features(h), (54, 161))
    assert np.array_equal(f2[:4], first_rows)            # context = last 4 rows of the previous chunk
    assert L.orc_nsnet2_denoise(h, po.f32p(out), 100, po.f32p(out)) != 0   # InvalidInputLength (NSNet2.zig:167-169)
    L.orc_nsnet2_free(h)
