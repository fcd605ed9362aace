"""Stage times of the reference-legal shape (48 kHz x 2 channels), device-resident: python tools/bench48.py [streams] [hours]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import fvad_b200
fv = fvad_b200.load()
S = int(sys.argv[1]) if len(sys.argv) > 1 else 20
hours = float(sys.argv[2]) if len(sys.argv) > 2 else 0.25
wd = fv.weights.synthetic(seed=1234)
x16, _ = fv.synth.onboard_batch_torch([1000 + s for s in range(S)], hours * 3600.0, 16000, "cuda:0")
nxt = torch.cat([x16[:, 1:], x16[:, -1:]], dim=1)
x = torch.empty((S, 2, x16.shape[1] * 3), dtype=torch.float32, device="cuda:0")
x[:, 0, 0::3] = x16
x[:, 0, 1::3] = x16 + (nxt - x16) * (1.0 / 3.0)
x[:, 0, 2::3] = x16 + (nxt - x16) * (2.0 / 3.0)
x[:, 1] = 0.8 * x[:, 0] + 0.2 * torch.roll(x[:, 0], 1, dims=0)
n = x.shape[2]
gp = fv.BatchPipeline(S, 48000, 2, wd, nn_mode=fv.capi.NN_TC_F16, device=0)
ptrs = [x[s, c].data_ptr() for s in range(S) for c in range(2)]
for _ in range(3):
    gp.push_pointers(ptrs, n, on_device=True)
acc = {}
for _ in range(3):
    gp.push_pointers(ptrs, n, on_device=True)
    st = gp.stats()
    for k in ("last_push_device_ms", "last_stft_ms", "last_nn_ms", "last_istft_ms", "last_vadfft_ms", "last_scan_ms"):
        acc[k] = acc.get(k, 0.0) + st[k] / 3
audio = S * (n // 24000) * 24000 / 48000
print(json.dumps({"streams": S, "hours": hours, "stream_audio_sec_per_sec": audio / (acc["last_push_device_ms"] * 1e-3),
                  **{k: round(v, 3) for k, v in acc.items()}}))
