set -x
timeout 600 python -m pytest tests/test_gpu_midsize.py -x -q > gpurun_out/c2_pytest_midsize.log 2>&1; echo "rc=$?" >> gpurun_out/c2_pytest_midsize.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/c2_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c2_pytest_all.log
timeout 600 python bench.py > gpurun_out/c2_bench_cfg2.json 2> gpurun_out/c2_bench_cfg2.err
timeout 600 python bench.py --asm-variant 2 --no-cpu-baseline > gpurun_out/c2_bench_cfg2_scatter.json 2> gpurun_out/c2_bench_cfg2_scatter.err
timeout 300 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c2_plain_128.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_assemble_rows -s 1 -c 1 -o gpurun_out/c2_rows_128 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c2_ncu_rows.log 2>&1
