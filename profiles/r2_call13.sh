set -x
timeout 300 python -m pytest tests/test_gpu_midsize.py tests/test_gpu_reentrancy.py -q -x > gpurun_out/c13_pytest_quick.log 2>&1; rc=$?; echo "rc=$rc" >> gpurun_out/c13_pytest_quick.log
FB2_SPMV_SCHED=1 timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c13_spmv_sweep_sched1_q2_256.txt 2>&1
FB2_SPMV_SCHED=0 timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c13_spmv_sweep_sched0_q2_256.txt 2>&1
FB2_SPMV_SCHED=1 timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c13_spmv_sweep_sched1_q3_128.txt 2>&1
FB2_SPMV_CONSUMERS=128 timeout 300 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c13_plain_128.log 2>&1 && \
FB2_SPMV_CONSUMERS=128 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_spmv_stream" -s 30 -c 1 -o gpurun_out/c13_spmv_sched_nc128_128 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c13_ncu.log 2>&1
