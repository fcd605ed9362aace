"""A small run of every kernel on the hot path (both sample rates, tensor-core mode, alt machines, evaluator) for
compute-sanitizer: python tools/sanitize_small.py"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import fvad_b200
fv = fvad_b200.load()
wd = fv.weights.synthetic(seed=1234)
for sr, nch, S in ((16000, 1, 3), (48000, 2, 2)):
    xs = []
    for s in range(S):
        x, _ = fv.synth.onboard_stream(10 + s, 6.2, 16000, nch)
        xs.append(fv.synth.expand_16k_to_48k(x) if sr == 48000 else x)
    x = np.stack(xs).reshape(S * nch, -1)
    gp = fv.BatchPipeline(S, sr, nch, wd, nn_mode=fv.capi.NN_TC_F16,
                          alt_vad_machine_configs=[fv.vad_machine_config(speech_threshold_factor=4.0)])
    gp.push(x[:, : x.shape[1] // 2 + 123])
    gp.push(x[:, x.shape[1] // 2 + 123:])
    print(sr, nch, [len(gp.segments(s)) for s in range(S)], gp.stats()["kernel_launches"])
    gp.close()
print("ok")
