# Final multi-GPU evidence on N GPUs of one box: NCCL / peer-memory parity tests, bench lines (plain, with the per-rank timeline,
# and with the peer-memory all-reduce switched off for the A/B).   usage: bash profiles/r2_final_par.sh N [short]
# short: the parity tests of tests/test_gpu_par.py and ONE bench line (with the per-rank timeline) only
N=$1
if [ "$2" = "short" ]; then
set -x
timeout 900 python -m pytest tests/test_gpu_par.py -m gpu -q > gpurun_out/fp_pytest_par_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/fp_pytest_par_n$N.log
tail -3 gpurun_out/fp_pytest_par_n$N.log
FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus $N > gpurun_out/fp_bench_n${N}_timeline.json 2> gpurun_out/fp_bench_n${N}_timeline.err
tail -c 1500 gpurun_out/fp_bench_n${N}_timeline.json
exit 0
fi
set -x
timeout 900 python -m pytest tests/test_gpu_par.py tests/test_gpu_boundary.py -m gpu -q > gpurun_out/fp_pytest_par_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/fp_pytest_par_n$N.log
tail -3 gpurun_out/fp_pytest_par_n$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus $N > gpurun_out/fp_bench_n$N.json 2> gpurun_out/fp_bench_n$N.err
FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus $N --no-e2e-assembly > gpurun_out/fp_bench_n${N}_timeline.json 2> gpurun_out/fp_bench_n${N}_timeline.err
FB2_PEER_ALLREDUCE=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29813 bench.py --gpus $N --no-e2e-assembly > gpurun_out/fp_bench_n${N}_nccl_allreduce.json 2> gpurun_out/fp_bench_n${N}_nccl_allreduce.err
python - <<PY
import json
for f in ("fp_bench_n$N", "fp_bench_n${N}_timeline", "fp_bench_n${N}_nccl_allreduce"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, "value %.0f" % d["value"], "ms_per_step %.2f" % d["ms_per_step"], "asm %.2f" % d["assembly_ms"], "pcg/it %.3f" % d["pcg_ms_per_iteration"],
              "spmv %.3f" % d["spmv_ms"], "e2e %.0f" % d["e2e"]["value"], "peer", d.get("peer_allreduce"), d["checksum"], d["clocks"])
    except Exception as e:
        print(f, "failed", e)
PY
