#!/bin/bash
# cfg4 step time under the knobs named on the command line: scripts/cfg4_sweep.sh "A=1 B=2" "A=3" ...
for envs in "$@"; do
  env $envs python bench.py --config cfg4 --batch 128 --n 512 --no-legs --no-cpu-baseline --steps 10 --warmup 3 > gpurun_out/cfg4_sweep.log 2>&1
  python - "$envs" <<PY
import json, sys
try:
    d = json.loads(open("gpurun_out/cfg4_sweep.log").read().strip().splitlines()[-1])
    print(sys.argv[1], "| ms", round(d["ms_per_step"], 4), "chain stage", round(d["roofline"]["stage_ms"], 4), "loss", d["e2e"]["loss_last"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
