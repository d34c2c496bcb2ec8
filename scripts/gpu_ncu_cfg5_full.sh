#!/bin/bash
# ncu --set full of the CTA-pair fp32x3 contraction inside the cfg5 step (eager launches; forward EPI 1 = scale, backward = row sums)
cd /root/repo
B="python bench.py --no-legs --no-cpu-baseline --n 2048 --steps 1 --warmup 3"
$B > gpurun_out/plain5.log 2>&1 || exit 1
DINO_B200_NO_GRAPH=1 timeout 1200 ncu --set full --clock-control none --kernel-name-base demangled \
  -k 'regex:gemm_tc2_kernel<\(bool\)1, \(int\)1' --launch-skip 48 -c 8 -o gpurun_out/r02d_cfg5_pair_full $B > gpurun_out/ncu5f.log 2>&1
echo rc $?
ls -la gpurun_out/r02d_cfg5_pair_full.ncu-rep
