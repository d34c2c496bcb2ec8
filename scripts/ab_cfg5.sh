#!/bin/bash
# A/B on one box: the worktree _ab_old (an earlier commit, built in place) against the current tree, alternating
cd /root/repo
for rep in 1 2; do
  for tree in _ab_old .; do
    ( cd $tree && python bench.py --no-legs --no-cpu-baseline --steps 20 --warmup 5 > /root/repo/gpurun_out/ab.log 2>&1
      python - "$tree" <<PY
import json, sys
try:
    d = json.loads(open("/root/repo/gpurun_out/ab.log").read().strip().splitlines()[-1])
    print(sys.argv[1], "| ms", round(d["ms_per_step"], 4), "chain stage", round(d["roofline"]["stage_ms"], 4), "frac", round(d["roofline"]["frac"], 4), "clk", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
    )
  done
done
