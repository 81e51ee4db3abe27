set -x
timeout 900 python -m pytest tests/test_gpu_par.py -q > gpurun_out/c11_pytest_par_n8.log 2>&1; echo "rc=$?" >> gpurun_out/c11_pytest_par_n8.log
FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus 8 > gpurun_out/c11_bench_n8.json 2> gpurun_out/c11_bench_n8.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus 8 --no-e2e-assembly > gpurun_out/c11_bench_n8_notimeline.json 2> gpurun_out/c11_bench_n8_notimeline.err
