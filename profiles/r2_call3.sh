set -x
# SpMV pattern classes first, small, with a short leash (a mismatch of the expected bytes would spin forever)
timeout 120 python -m pytest tests/test_gpu_midsize.py -x -q -k "spmv_variants or serial_entrywise" > gpurun_out/c3_pytest_spmv.log 2>&1; rc=$?; echo "rc=$rc" >> gpurun_out/c3_pytest_spmv.log
if [ $rc -ne 0 ]; then export FB2_SPMV_CLASSES=0; echo "classes disabled for the rest of the call" >> gpurun_out/c3_pytest_spmv.log; fi
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/c3_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c3_pytest_all.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/c3_bench_cfg2.json 2> gpurun_out/c3_bench_cfg2.err
timeout 600 python bench.py --no-cpu-baseline --spmv-variant 3 --no-e2e-assembly > gpurun_out/c3_bench_cfg2_noclasses.json 2> gpurun_out/c3_bench_cfg2_noclasses.err
timeout 600 python bench.py --no-cpu-baseline --order 3 --no-e2e-assembly > gpurun_out/c3_bench_cfg5.json 2> gpurun_out/c3_bench_cfg5.err
timeout 600 python bench.py --no-cpu-baseline --kind vector_laplacian > gpurun_out/c3_bench_cfg4.json 2> gpurun_out/c3_bench_cfg4.err
timeout 300 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c3_plain_128.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_assemble_rows|k_spmv_stream" -s 2 -c 4 -o gpurun_out/c3_rows_spmv_128 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c3_ncu.log 2>&1
