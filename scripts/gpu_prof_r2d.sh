#!/bin/bash
# round-2 final measurements: every BASELINE configuration once (fp32x3 and tf32), then the launch lists of cfg5 and cfg4
cd /root/repo
for cfg in cfg1 cfg2 cfg3 cfg3b cfg4 cfg5; do for p in fp32x3 tf32; do
  python bench.py --config $cfg --precision $p --no-legs --no-cpu-baseline --steps 20 --warmup 5 > gpurun_out/r2d_${cfg}_${p}.log 2>&1
  python - "$cfg $p" <<PY
import json, sys
try:
    d = json.loads(open("gpurun_out/r2d_${cfg}_${p}.log").read().strip().splitlines()[-1])
    print(sys.argv[1], "| B", d["config"]["per_gpu_batch"], "| ms", round(d["ms_per_step"], 4), "| samples/s", round(d["value"]), "| e2e", round(d["e2e"]["value"]),
          "| step TF/s", round(d["roofline"]["step"]["achieved_tflops"], 1), "| stage frac", round(d["roofline"]["frac"], 3), "| launches", d["gpu_launches"])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
done; done
python bench.py --no-legs --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/plain5.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02d_ncu_launches_cfg5_fp32x3.csv python bench.py --no-legs --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/ncu5.log 2>&1
echo ncu5 rc $?
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02d_ncu_launches_cfg4_fp32x3.csv python bench.py --config cfg4 --batch 128 --n 512 --no-legs --no-cpu-baseline --steps 2 --warmup 3 > gpurun_out/ncu4.log 2>&1
echo ncu4 rc $?
