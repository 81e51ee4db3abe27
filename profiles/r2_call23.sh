# repeatability of the cold start / class set-up at config 2
for k in 1 2 3; do
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c23_bench_$k.json 2> gpurun_out/c23_bench_$k.err
python - <<PY
import json
d=json.loads(open("gpurun_out/c23_bench_$k.json").read().strip().splitlines()[-1])
print("run $k", d["setup_ms"], "cold", d["cold_start_ms"], "spmv GB/s", d["value"], "clocks", d["clocks"])
PY
done
