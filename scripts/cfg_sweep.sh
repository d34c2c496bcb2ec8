#!/bin/bash
# step time of one configuration under the knobs named on the command line: scripts/cfg_sweep.sh cfg3 "A=1 B=2" "A=3" ...
cfg=$1; shift
for envs in "$@"; do
  env $envs python bench.py --config $cfg --no-legs --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/cfg_sweep.log 2>&1
  python - "$cfg $envs" <<PY
import json, sys
try:
    d = json.loads(open("gpurun_out/cfg_sweep.log").read().strip().splitlines()[-1])
    print(sys.argv[1], "| ms", round(d["ms_per_step"], 4), "chain stage", round(d["roofline"]["stage_ms"], 4), "loss", d["e2e"]["loss_last"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done
