"""Print the tcgen05 path's gain error against the golden fixtures (abs max, and relative to max)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fvad_b200
fv = fvad_b200.load()
for name, sr, nch in [("pipeline_16k_mono.npz", 16000, 1), ("pipeline_48k_stereo.npz", 48000, 2)]:
    g = np.load(os.path.join(ROOT, "tests", "golden", name))
    x = g["pcm16"].astype(np.float32) / np.float32(32768.0)
    if sr == 48000:
        x = fv.synth.expand_16k_to_48k(x)
    wd = fv.weights.synthetic(seed=int(g["weights_seed"]))
    for mode, nm in [(fv.capi.NN_F32, "f32"), (fv.capi.NN_TC_F16, "f16")]:
        gp = fv.BatchPipeline(1, sr, nch, wd, nn_mode=mode, keep_taps=True)
        gp.push(x)
        gain = gp.tap(fv.capi.TAP_GAINS)[0][g["frame_rows"]]
        e = np.abs(gain - g["gain"])
        met = (gp.tap(fv.capi.TAP_MET)[0] == g["met"]).mean()
        print(f"{name} {nm}: gains abs max {e.max():.3e} rel-to-max {e.max() / np.abs(g['gain']).max():.3e} rms {np.sqrt((e**2).mean()):.3e} decisions {met:.5f}")
        gp.close()
