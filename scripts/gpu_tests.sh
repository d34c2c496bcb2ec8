# full -m gpu suite in chunks (each under its own timeout: a dead-locked mbarrier protocol must not hang the box)
cd /root/repo
export PYTHONPATH=/root/repo
for f in test_gpu_tc test_gpu_parity test_gpu_shapes test_gpu_properties test_gpu_driver test_gpu_inference test_gpu_data test_gpu_multi test_zz_no_alloc; do
  timeout 1200 python -m pytest tests/$f.py -q -m gpu -x 2>&1 | tail -${TAILN:-12} > gpurun_out/t_$f.log; echo "$f exit ${PIPESTATUS[0]}" >> gpurun_out/t_$f.log
  cat gpurun_out/t_$f.log
done
