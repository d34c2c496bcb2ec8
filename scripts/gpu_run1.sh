cd /root/repo
export PYTHONPATH=/root/repo
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests/test_gpu_tc.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/t_tc.log; echo "tc exit $?" >> gpurun_out/t_tc.log
DINO_TC_VARIANT=1 timeout 600 python -m pytest tests/test_gpu_tc.py -x -q -m gpu -k "gemm_matches" 2>&1 | tail -8 > gpurun_out/t_tc_v1.log
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -25 > gpurun_out/t_parity.log
timeout 1500 python -m pytest tests/test_gpu_shapes.py -q -m gpu 2>&1 | tail -60 > gpurun_out/t_shapes.log
cat gpurun_out/t_tc.log gpurun_out/t_tc_v1.log gpurun_out/t_parity.log gpurun_out/t_shapes.log
