# Final round-2 evidence on ONE B200: GPU tests, bench lines of configs 2 / 4 / 5, reference arm, ncu launch list of the bench
# command (cold-cache, serialised: compare SHARES), DRAM bytes of the SpMV launches at config 2 (-> profiles/spmv_traffic.json).
set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/f_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/f_pytest_all.log
timeout 600 python bench.py > gpurun_out/f_bench_cfg2.json 2> gpurun_out/f_bench_cfg2.err
timeout 600 python bench.py --kind vector_laplacian --no-cpu-baseline > gpurun_out/f_bench_cfg4_veclap.json 2> gpurun_out/f_bench_cfg4_veclap.err
timeout 600 python bench.py --kind elasticity --no-cpu-baseline > gpurun_out/f_bench_cfg4_elast.json 2> gpurun_out/f_bench_cfg4_elast.err
timeout 600 python bench.py --order 3 --no-cpu-baseline > gpurun_out/f_bench_cfg5.json 2> gpurun_out/f_bench_cfg5.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/f_bench_ref.json 2> gpurun_out/f_bench_ref.err
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/f_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/f_launches_q2_256.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/f_ncu_launches.log 2>&1 && \
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"k_spmv_stream|k_assemble_rows" -s 2 -c 6 --csv --log-file gpurun_out/f_dram_q2_256.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/f_ncu_dram.log 2>&1
