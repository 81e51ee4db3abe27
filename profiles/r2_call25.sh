# ncu of the SpMV launches of one sub-domain (8 co-resident parts of config 3 on one GPU): what differs from the serial 128^3 matrix
set -x
timeout 900 ncu --set full --clock-control none -k regex:k_spmv_stream -s 40 -c 4 -o gpurun_out/c25_spmv_part python bench.py --subdomains 8 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c25_ncu.log 2>&1
tail -2 gpurun_out/c25_ncu.log
