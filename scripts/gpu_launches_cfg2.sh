#!/bin/bash
# launch list of the cfg2 step (fused chain kernel), eager launches
cd /root/repo
B="python bench.py --config cfg2 --no-legs --no-cpu-baseline --steps 3 --warmup 3"
$B > gpurun_out/plain2.log 2>&1 || exit 1
DINO_B200_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02d_ncu_launches_cfg2_fp32x3.csv $B > gpurun_out/ncu2.log 2>&1
echo rc $?
