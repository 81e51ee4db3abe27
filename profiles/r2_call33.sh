# SpMV: bank-preserving x-window layout, parity (bitwise variants test) then A/B at config 2 / config 5 / vector sizes
set -x
timeout 600 python -m pytest tests/test_gpu_midsize.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/c33_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c33_pytest.log
tail -3 gpurun_out/c33_pytest.log
timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c33_sweep_q2_256.txt 2>&1
cat gpurun_out/c33_sweep_q2_256.txt
timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c33_sweep_q3_128.txt 2>&1
cat gpurun_out/c33_sweep_q3_128.txt
