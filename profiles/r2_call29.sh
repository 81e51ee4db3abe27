# row-owner assembly: transposed plan table + swizzled chunk image; parity, then A/B/A/B of the swizzle at config 2
set -x
timeout 900 python -m pytest tests/test_gpu_midsize.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -q -x > gpurun_out/c29_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c29_pytest.log
tail -3 gpurun_out/c29_pytest.log
for sw in 1 0 1 0; do
FB2_ASM_SWIZZLE=$sw timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c29_bench_swz$sw.json 2> gpurun_out/c29_bench_swz$sw.err
python - <<PY
import json
d=json.loads(open("gpurun_out/c29_bench_swz$sw.json").read().strip().splitlines()[-1])
print("swizzle=$sw asm %.2f ms %.0f MDOF/s spmv %.3f"%(d["assembly_ms"], d["assembly_mdof_per_s"], d["spmv_ms"]), d["checksum"], d["clocks"]["sm_mhz"])
PY
done
