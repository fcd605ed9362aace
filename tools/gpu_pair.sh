#!/bin/bash
# pair-GEMM bring-up: parity on the golden fixtures under a timeout, then a same-box A/B
mkdir -p gpurun_out
FVAD_GEMM_PAIR=1 timeout -s KILL 120 python tools/gain_err.py 2>&1 | tail -6
echo "gain_err rc=$?"
timeout -s KILL 300 bash tools/gpu_ab.sh FVAD_GEMM_PAIR=0 FVAD_GEMM_PAIR=1 2>&1 | tail -4
for v in 0 1; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_FVAD_GEMM_PAIR=$v.json")); print("pair=$v", {k:round(x,2) for k,x in d["nn_layer_ms_per_step"].items()})
except Exception as e: print("pair=$v", e); print(open("gpurun_out/ab_FVAD_GEMM_PAIR=$v.err").read()[-1500:])
PY
done
