# CSR graph builder with learned permutations (k_row_fill_perm), scatter-map set-up trims: parity, then set-up times at configs 2 / 4 / 5
set -x
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/c21_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c21_pytest.log
tail -3 gpurun_out/c21_pytest.log
for v in 2 1; do
FB2_ROW_FILL=$v timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c21_bench_cfg2_fill$v.json 2> gpurun_out/c21_bench_cfg2_fill$v.err
python - <<PY
import json
d=json.loads(open("gpurun_out/c21_bench_cfg2_fill$v.json").read().strip().splitlines()[-1])
print("fill=$v", d["setup_ms"], "cold", d["cold_start_ms"], "spmv GB/s", d["value"], "asm", d["assembly_ms"], d["checksum"])
PY
done
timeout 600 python bench.py --kind vector_laplacian --steps 2 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c21_bench_cfg4.json 2> gpurun_out/c21_bench_cfg4.err
timeout 600 python bench.py --order 3 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c21_bench_cfg5.json 2> gpurun_out/c21_bench_cfg5.err
python - <<PY
import json
for f in ("cfg4","cfg5"):
    d=json.loads(open("gpurun_out/c21_bench_%s.json"%f).read().strip().splitlines()[-1])
    print(f, d["setup_ms"], "cold", d["cold_start_ms"], "spmv GB/s", d["value"], "asm", d["assembly_ms"], d["checksum"])
PY
