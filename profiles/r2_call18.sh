# transposed row map + staggered chunks in k_spmv_stream: parity first, then A/B at config 2 and config 5 sizes, then ncu
set -x
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_midsize.py tests/test_gpu_reentrancy.py tests/test_gpu_boundary.py -m gpu -q -x > gpurun_out/c18_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c18_pytest.log
tail -5 gpurun_out/c18_pytest.log
timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c18_sweep_q2_256.txt 2>&1
cat gpurun_out/c18_sweep_q2_256.txt
timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c18_sweep_q3_128.txt 2>&1
cat gpurun_out/c18_sweep_q3_128.txt
timeout 300 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c18_plain.log 2>&1 && \
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_spmv_stream -s 10 -c 1 -o gpurun_out/c18_spmv_tr_128 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c18_ncu.log 2>&1
tail -3 gpurun_out/c18_ncu.log
