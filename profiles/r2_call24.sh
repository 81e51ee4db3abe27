# private tile descriptors: parity, then A/B at config 2 (serial) and on the 8 co-resident sub-domains of config 3 (one GPU)
set -x
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/c24_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/c24_pytest.log
tail -3 gpurun_out/c24_pytest.log
for pv in 1 0; do
FB2_SPMV_PRIVATE=$pv timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c24_bench_cfg2_priv$pv.json 2> gpurun_out/c24_bench_cfg2_priv$pv.err
FB2_SPMV_PRIVATE=$pv timeout 600 python bench.py --subdomains 8 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c24_bench_sub8_priv$pv.json 2> gpurun_out/c24_bench_sub8_priv$pv.err
python - <<PY
import json
for f in ("cfg2","sub8"):
    try:
        d=json.loads(open("gpurun_out/c24_bench_%s_priv$pv.json"%f).read().strip().splitlines()[-1])
        print(f, "private=$pv value %.0f spmv %.3f pcg/it %.3f"%(d["value"], d["spmv_ms"], d["pcg_ms_per_iteration"]), d["spmv_tiles"], d["checksum"], d["clocks"]["sm_mhz"])
    except Exception as e:
        print(f, "failed", e, open("gpurun_out/c24_bench_%s_priv$pv.err"%f).read()[-600:])
PY
done
