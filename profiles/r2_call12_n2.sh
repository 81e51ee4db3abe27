set -x
timeout 900 python -m pytest tests/test_gpu_par.py -q > gpurun_out/c12_pytest_par_n2.log 2>&1; echo "rc=$?" >> gpurun_out/c12_pytest_par_n2.log
for pa in 1 0; do
FB2_PEER_ALLREDUCE=$pa FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2971$pa bench.py --gpus 2 --cells 128 --no-e2e-assembly > gpurun_out/c12_bench_n2_128_peer$pa.json 2> gpurun_out/c12_bench_n2_128_peer$pa.err
done
FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 > gpurun_out/c12_bench_n2.json 2> gpurun_out/c12_bench_n2.err
