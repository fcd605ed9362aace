"""GPU debug: tensor-core NN backend vs the f32 SIMT backend and the oracle on a short batch."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fvad_b200
fv = fvad_b200.load()
wd = fv.weights.synthetic(seed=1234)
S = int(sys.argv[1]) if len(sys.argv) > 1 else 3
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 30.0
xs = np.concatenate([fv.synth.onboard_stream(100 + s, secs, 16000, 1)[0] for s in range(S)], axis=0)
res = {}
for name, mode in (("f32", fv.capi.NN_F32), ("f16", fv.capi.NN_TC_F16)):
    gp = fv.BatchPipeline(S, 16000, 1, wd, nn_mode=mode, keep_taps=True)
    t = time.time(); gp.push(xs); dt = time.time() - t
    res[name] = (gp.tap(fv.capi.TAP_GAINS), gp.tap(fv.capi.TAP_BAND), gp.tap(fv.capi.TAP_MET), [gp.segments(s) for s in range(S)])
    print(name, "push %.3fs" % dt, gp.stats())
    gp.close()
g0, g1 = res["f32"][0], res["f16"][0]
print("gains: max abs diff %.3e  mean abs diff %.3e  nan %d" % (np.abs(g0 - g1).max(), np.abs(g0 - g1).mean(), np.isnan(g1).sum()))
d = np.abs(g0 - g1).reshape(-1, 161).max(axis=1)
print("worst rows:", np.argsort(d)[-5:], d[np.argsort(d)[-5:]], "rows>1e-2:", (d > 1e-2).sum(), "of", d.size)
b0, b1 = res["f32"][1], res["f16"][1]
print("band rel diff %.3e" % (np.abs(b0 - b1).max() / np.abs(b0).max()))
print("decisions agree %.5f" % (res["f32"][2] == res["f16"][2]).mean())
print("segments f32:", res["f32"][3][0][:3], "f16:", res["f16"][3][0][:3])
