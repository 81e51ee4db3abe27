set -x
python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
python bench.py > gpurun_out/c1_bench_cfg2.json 2> gpurun_out/c1_bench_cfg2.err
python bench.py --kind vector_laplacian --no-cpu-baseline > gpurun_out/c1_bench_cfg4_veclap.json 2> gpurun_out/c1_bench_cfg4_veclap.err
python bench.py --kind elasticity --no-cpu-baseline > gpurun_out/c1_bench_cfg4_elast.json 2> gpurun_out/c1_bench_cfg4_elast.err
python bench.py --order 3 --no-cpu-baseline > gpurun_out/c1_bench_cfg5.json 2> gpurun_out/c1_bench_cfg5.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/c1_bench_ref.json 2> gpurun_out/c1_bench_ref.err
python bench.py --order 3 --cells 64 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c1_plain_q3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_assemble_fast -s 1 -c 1 -o gpurun_out/c1_q3_asm python bench.py --order 3 --cells 64 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c1_ncu_q3.log 2>&1
nvidia-smi > gpurun_out/c1_smi.txt
