#!/bin/bash
# A/B on one box; prints per-layer NN ms
for v in "$@"; do
  env $v python bench.py --steps 2 --warmup 2 --hours 1 --no-cpu-baseline --no-48k --e2e-steps 1 > gpurun_out/ab2.json 2>gpurun_out/ab2.err
  python - <<PY
import json
d=json.load(open("gpurun_out/ab2.json"))
print("$v", "value=%.0f"%d["value"], "nn=%.1f"%d["stage_ms_per_step"]["nsnet2"], {k:round(x,2) for k,x in d["nn_layer_ms_per_step"].items()}, "mhz", d["clocks"]["sm_mhz"], "W", d["clocks"].get("power_w_median"))
PY
done
