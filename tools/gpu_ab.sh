#!/bin/bash
# A/B: same box, env var toggles; prints nsnet2 stage ms
mkdir -p gpurun_out
for v in "$@"; do
  env $v python bench.py --steps 3 --warmup 2 --hours 1 --no-cpu-baseline --no-48k --e2e-steps 1 > gpurun_out/ab_$v.json 2>gpurun_out/ab_$v.err
  python - <<PY
import json
d=json.load(open("gpurun_out/ab_$v.json"))
print("$v", "value=%.0f"%d["value"], {k:round(x,2) for k,x in d["stage_ms_per_step"].items()}, d["clocks"])
PY
done
