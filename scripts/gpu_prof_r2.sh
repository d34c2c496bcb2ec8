cd /root/repo
B="python bench.py --n 2048 --no-legs --no-cpu-baseline --steps 2 --warmup 3"
$B > gpurun_out/plain_r2b.log 2>&1 &&
DINO_B200_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_r2b.csv $B > gpurun_out/ncu_r2b.log 2>&1
echo launches rc $?
$B > gpurun_out/plain_r2b.log 2>&1 &&
DINO_B200_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 230 -c 4 -o gpurun_out/prof_gemm_x3_r2b $B > gpurun_out/ncu_r2b_full.log 2>&1
echo full rc $?
