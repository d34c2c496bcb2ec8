#!/bin/bash
# ncu --set full of the two gathered contractions of the cfg4 step (J W_1 and J^T (c Q)); eager launches
cd /root/repo
B="python bench.py --config cfg4 --batch 128 --n 256 --no-legs --no-cpu-baseline --steps 1 --warmup 3"
$B > gpurun_out/plain4.log 2>&1 || exit 1
DINO_B200_NO_GRAPH=1 timeout 1200 ncu --set full --clock-control none --kernel-name-base demangled \
  -k 'regex:gemm_tc_kernel<\(int\)128, \(int\)32, \(int\)3, \(bool\)1, \(int\)1, \(int\)4, \(int\)8>|wgrad_tc_kernel<\(int\)128, \(int\)32, \(int\)3, \(bool\)1>' \
  --launch-skip 8 -c 16 -o gpurun_out/r02d_cfg4_full $B > gpurun_out/ncu4f.log 2>&1
echo rc $?
ls -la gpurun_out/r02d_cfg4_full.ncu-rep
