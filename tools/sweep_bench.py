"""BASELINE config 5: VAD threshold/hysteresis sweep -- 256 alternative machine configs over 20 streams, scored
by the on-device Evaluator.  Prints one JSON line: configs x streams x audio-seconds scanned per second."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import fvad_b200

ap = argparse.ArgumentParser()
ap.add_argument("--streams", type=int, default=20)
ap.add_argument("--hours", type=float, default=2.0)
ap.add_argument("--configs", type=int, default=256)
args = ap.parse_args()
fv = fvad_b200.load()
wd = fv.weights.synthetic(seed=1234)
S, sr = args.streams, 16000
x, labels = fv.synth.onboard_batch_torch([1000 + s for s in range(S)], args.hours * 3600.0, sr, "cuda:0")
n = x.shape[1]
ptrs = [x[s].data_ptr() for s in range(S)]
rng = np.random.default_rng(0)
alts = [fv.vad_machine_config(speech_threshold_factor=float(f), max_speech_gap_sec=float(g), min_consecutive_sec_to_open=float(o))
        for f, g, o in zip(rng.uniform(2, 30, args.configs), rng.uniform(0.5, 4, args.configs), rng.uniform(0.1, 0.6, args.configs))]
res = {}
for name, a in (("main_only", None), ("sweep", alts)):
    gp = fv.BatchPipeline(S, sr, 1, wd, alt_vad_machine_configs=a, nn_mode=fv.capi.NN_TC_F16)
    gp.push_pointers(ptrs, n, on_device=True)   # warm-up
    t = []
    for _ in range(2):
        gp.push_pointers(ptrs, n, on_device=True)
        st = gp.stats()
        t.append((st["last_push_device_ms"], st["last_scan_ms"]))
    res[name] = (min(v[0] for v in t), min(v[1] for v in t))
    if a:
        refs = [[(float(np.float32(p)), float(np.float32(q))) for p, q in lab] for lab in labels]
        cfg = fv.eval_config(extrude_start=2.0, extrude_end=2.0, fill_gaps=2.0)
        t0 = time.perf_counter(); table = gp.evaluate(refs, cfg); ev_ms = 1e3 * (time.perf_counter() - t0)
        f = np.array([[table[s][m]["f_score"] for m in range(1 + len(a))] for s in range(S)])
        best = int(np.nanargmax(np.nanmean(f, axis=0)))
        nseg = sum(len(gp.segments(s, m)) for s in range(S) for m in range(0, 1 + len(a), 16))
    gp.close()
audio = S * (n // 8000) * 8000 / sr
extra = res["sweep"][0] - res["main_only"][0]
print(json.dumps({"workload": f"config 5: {args.configs} alt VAD machine configs x {S} streams x {args.hours:g} h, full pipeline once + scan per config + device Evaluator",
                  "push_ms_main_only": res["main_only"][0], "push_ms_with_sweep": res["sweep"][0], "scan_ms_main_only": res["main_only"][1],
                  "scan_ms_sweep": res["sweep"][1], "evaluate_ms": ev_ms, "lists_scored": S * (1 + args.configs),
                  "config_audio_sec_per_sec": args.configs * audio / (max(extra, 1e-3) * 1e-3 + ev_ms * 1e-3),
                  "best_config_index": best, "segments_sampled": nseg}))
