# fused PCG set-up (zero initial solution): full GPU suite + the default bench line
set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c27_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c27_pytest_all.log
tail -4 gpurun_out/c27_pytest_all.log
timeout 600 python bench.py > gpurun_out/c27_bench_cfg2.json 2> gpurun_out/c27_bench_cfg2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/c27_bench_cfg2.json").read().strip().splitlines()[-1])
print("value %.0f frac %.3f asm %.2f pcg/it %.3f spmv %.3f e2e %.0f cold %.0f"%(d["value"], d["roofline"]["frac"], d["assembly_ms"], d["pcg_ms_per_iteration"], d["spmv_ms"], d["e2e"]["value"], d["cold_start_ms"]), d["checksum"], d["clocks"], d["roofline"].get("traffic_frac_of_peak"))
PY
