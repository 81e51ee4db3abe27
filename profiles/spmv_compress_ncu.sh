set -x
B="python bench.py --cells 128 --steps 1 --warmup 1 --cg-iters 3 --no-cpu-baseline"
$B --spmv-variant 2 > gpurun_out/comp_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_spmv_stream -s 4 -c 1 -f -o gpurun_out/spmv_comp_q2_128 $B --spmv-variant 2 > gpurun_out/comp_ncu.log 2>&1
