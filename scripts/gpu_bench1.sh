cd /root/repo
B="python bench.py --no-fp32-leg --no-cpu-baseline --random-data"
for cfg in cfg5 cfg3; do for p in fp32x3 tf32; do
  $B --config $cfg --precision $p --n 2048 --steps 10 --warmup 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$cfg $p', 'ms/step', round(d['ms_per_step'],3), 'samples/s', round(d['value']), 'TF/s', round(d['roofline']['step']['achieved_tflops'],1), 'e2e', round(d['e2e']['value']), 'graph', d['config']['cuda_graph_step'])"
done; done
for p in fp32x3 tf32; do
$B --config cfg5 --precision $p --n 2048 --steps 2 --warmup 3 > gpurun_out/plain_cfg5.log 2>&1 &&
DINO_B200_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_cfg5_$p.csv $B --config cfg5 --precision $p --n 2048 --steps 2 --warmup 3 > gpurun_out/ncu_cfg5.log 2>&1
echo ncu rc $?
done
