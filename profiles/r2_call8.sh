set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/c8_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c8_pytest_all.log
timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c8_spmv_sweep_q2_256.txt 2>&1
timeout 600 python bench.py > gpurun_out/c8_bench_cfg2.json 2> gpurun_out/c8_bench_cfg2.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/c8_bench_ref.json 2> gpurun_out/c8_bench_ref.err
