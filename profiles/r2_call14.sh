set -x
timeout 300 python -m pytest tests/test_gpu_midsize.py tests/test_gpu_parity.py -q -x -k "spmv or midsize_serial" > gpurun_out/c14_pytest_quick.log 2>&1; rc=$?; echo "rc=$rc" >> gpurun_out/c14_pytest_quick.log
FB2_SPMV_SCHED=1 timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c14_spmv_sweep_sched1_q2_256.txt 2>&1
FB2_SPMV_SCHED=1 timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c14_spmv_sweep_sched1_q3_128.txt 2>&1
FB2_SPMV_SCHED=0 timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c14_spmv_sweep_sched0_q3_128.txt 2>&1
