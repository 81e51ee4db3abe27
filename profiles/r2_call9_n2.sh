set -x
timeout 900 python -m pytest tests/test_gpu_par.py -q > gpurun_out/c9_pytest_par_n2.log 2>&1; echo "rc=$?" >> gpurun_out/c9_pytest_par_n2.log
FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 > gpurun_out/c9_bench_n2.json 2> gpurun_out/c9_bench_n2.err
FB2_PAR_TIMELINE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 --subdomains 2 --no-e2e-assembly > gpurun_out/c9_bench_n2_sub2.json 2> gpurun_out/c9_bench_n2_sub2.err
