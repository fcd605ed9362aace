"""Generate tests/golden/pipeline_16k_mono.npz and pipeline_48k_stereo.npz: known-answer vectors for the
hot path, produced by the CPU oracle with its real FFTs routed through the REFERENCE's own kissfft
(oracle/_ref/libkissfft_ref.so, compiled unmodified from /root/reference/lib/kissfft).  Run in the build
container (where /root/reference exists):   python tools/make_golden.py
The NSNet2 weights are the seeded synthetic set (weights.synthetic(1234)): the shipped .onnx is absent
from the reference snapshot, so gains are pinned to the restated network, not to onnxruntime."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fvad_b200  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

fv = fvad_b200.load()


def make_input(seed, seconds, n_channels):
    """16 kHz, int16-quantised (exact in f32): low engine/broadband noise + two loud speech-like bursts."""
    rng = np.random.default_rng(seed)
    n = int(seconds * 16000)
    t = np.arange(n) / 16000.0
    chans = []
    burst = np.zeros(n)
    for (a, b) in ((1.2, 3.1), (6.4, 7.0)):
        m = (t >= a) & (t < b)
        env = 0.6 + 0.4 * np.sin(2 * np.pi * 4 * t[m])
        burst[m] = 0.25 * env * sum(np.sin(2 * np.pi * f * t[m] + p) for f, p in ((620, 0.3), (1180, 1.1), (1730, 2.0)))
    for c in range(n_channels):
        eng = 0.004 * sum(np.sin(2 * np.pi * 97 * h * t + h) / h for h in range(1, 9))
        x = (1.0 if c == 0 else 0.8) * eng + rng.normal(0, 0.0015, n) + burst
        chans.append(np.round(np.clip(x, -1, 1) * 32767).astype(np.int16))
    return np.stack(chans)


def run(name, sr, nch, seconds, seed):
    assert po.use_reference_kissfft(True), "oracle/_ref/libkissfft_ref.so missing: run `make -C oracle ref`"
    wd = fv.weights.synthetic(seed=1234)
    pcm16 = make_input(seed, seconds, nch)
    x = pcm16.astype(np.float32) / np.float32(32768.0)
    if sr == 48000:
        x = fv.synth.expand_16k_to_48k(x)
    alts = [dict(speech_threshold_factor=4.0), dict(speech_threshold_factor=16.0, max_speech_gap_sec=1.0)]
    p = po.Pipeline(wd, sr, nch, alts=[po.VadMachineConfig.default(**a) for a in alts], keep_denoised=True, keep_nn=True)
    p.push_samples(x)
    feat, gain, spec = p.nn_taps()
    band, ratio, met = p.frames()
    segs = [np.array(p.segments(m), dtype=np.float64).reshape(-1, 4) for m in range(3)]
    den = np.stack([p.denoised(c) for c in range(nch)])
    fr = np.arange(0, feat.shape[0], 13)
    out = dict(pcm16=pcm16, sample_rate=sr, weights_seed=1234, frame_rows=fr, spec=spec[fr], feat=feat[fr], gain=gain[fr],
               denoised_stride16=den[:, ::16].astype(np.float32), band=band, ratio=ratio, met=met,
               segs0=segs[0], segs1=segs[1], segs2=segs[2])
    path = os.path.join(ROOT, "tests", "golden", name)
    np.savez_compressed(path, **out)
    print(name, "frames", feat.shape[0], "vad frames", band.size, "met", int(met.sum()), "segments", [len(s) for s in segs],
          "bytes", os.path.getsize(path))
    po.use_reference_kissfft(False)


if __name__ == "__main__":
    run("pipeline_16k_mono.npz", 16000, 1, 10.0, 42)
    run("pipeline_48k_stereo.npz", 48000, 2, 10.0, 43)
