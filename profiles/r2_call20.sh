# SpMV: consumer passes (128 consumer threads over 32-row tiles), A/B against 256
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_midsize.py -m gpu -q -x > gpurun_out/c20_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c20_pytest.log
tail -3 gpurun_out/c20_pytest.log
timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c20_sweep_q2_256.txt 2>&1
cat gpurun_out/c20_sweep_q2_256.txt
timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c20_sweep_q3_128.txt 2>&1
cat gpurun_out/c20_sweep_q3_128.txt
