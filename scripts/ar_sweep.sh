#!/bin/bash
# all-reduce timings under NCCL knobs: scripts/ar_sweep.sh N "A=1" "B=2 C=3" ...   (log: gpurun_out/ar_sweep.log)
cd /root/repo
N=$1; shift
for envs in "$@"; do
  env $envs timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 scripts/ar_bench.py 2>&1 | grep -E "^n=|Error" | tee -a gpurun_out/ar_sweep.log
done
