# L2 hint default on: full GPU suite, default bench line, DRAM bytes of the SpMV launches
set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c31_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c31_pytest_all.log
tail -3 gpurun_out/c31_pytest_all.log
timeout 600 python bench.py > gpurun_out/c31_bench_cfg2.json 2> gpurun_out/c31_bench_cfg2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/c31_bench_cfg2.json").read().strip().splitlines()[-1])
print("value %.0f frac %.3f asm %.2f pcg/it %.3f spmv %.3f e2e %.0f cold %.0f"%(d["value"], d["roofline"]["frac"], d["assembly_ms"], d["pcg_ms_per_iteration"], d["spmv_ms"], d["e2e"]["value"], d["cold_start_ms"]), d["checksum"], d["clocks"])
PY
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"k_spmv_stream" -s 2 -c 6 --csv --log-file gpurun_out/c31_dram_q2_256.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c31_ncu_dram.log 2>&1
tail -2 gpurun_out/c31_ncu_dram.log | cut -c1-200
