# source-level ncu capture of the final SpMV kernel at Q2 128^3
set -x
timeout 900 ncu --set full --import-source on --clock-control none -k regex:k_spmv_stream -s 10 -c 1 -o gpurun_out/c32_spmv_128 python bench.py --cells 128 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c32_ncu.log 2>&1
tail -2 gpurun_out/c32_ncu.log | cut -c1-200
