#!/bin/bash
# launch list of the cfg5 step (eager launches, small resident dataset so that the list starts near the steps)
cd /root/repo
B="python bench.py --no-legs --no-cpu-baseline --n 2048 --steps 2 --warmup 3"
$B > gpurun_out/plain5.log 2>&1 || exit 1
DINO_B200_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02d_ncu_launches_cfg5_fp32x3.csv $B > gpurun_out/ncu5.log 2>&1
echo rc $?
