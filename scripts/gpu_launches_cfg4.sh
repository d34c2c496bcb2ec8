#!/bin/bash
# launch list of the cfg4 step (eager launches)
cd /root/repo
B="python bench.py --config cfg4 --batch 128 --n 512 --no-legs --no-cpu-baseline --steps 2 --warmup 3"
$B > gpurun_out/plain4.log 2>&1 || exit 1
DINO_B200_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02d_ncu_launches_cfg4_fp32x3.csv $B > gpurun_out/ncu4.log 2>&1
echo rc $?
