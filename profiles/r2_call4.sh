set -x
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/c4_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c4_pytest_all.log
timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c4_spmv_sweep_q2_256.txt 2>&1
timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c4_spmv_sweep_q3_128.txt 2>&1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/c4_bench_cfg2.json 2> gpurun_out/c4_bench_cfg2.err
timeout 300 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c4_plain_128.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_assemble_rows" -s 1 -c 1 -o gpurun_out/c4_rows_128 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c4_ncu.log 2>&1
