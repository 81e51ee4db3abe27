set -x
B1="python bench.py --steps 1 --warmup 1 --cg-iters 5 --no-cpu-baseline"
B2="python bench.py --cells 128 --steps 2 --warmup 1 --cg-iters 10 --no-cpu-baseline"
B3="python bench.py --cells 128 --steps 1 --warmup 1 --cg-iters 3 --no-cpu-baseline"
$B1 > gpurun_out/r1f_plain256.log 2>&1 && ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_spmv_stream -s 4 -c 2 --csv --log-file gpurun_out/r1f_spmv256_dram.csv $B1 > gpurun_out/r1f_ncu256.log 2>&1
$B2 > gpurun_out/r1f_plain128.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1f_launches_q2_128.csv $B2 > gpurun_out/r1f_ncu128.log 2>&1
$B3 > gpurun_out/r1f_plain128b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"k_assemble_fast|k_spmv_stream|k_cg_update" -s 3 -c 5 -f -o gpurun_out/r1f_full_q2_128 $B3 > gpurun_out/r1f_ncufull.log 2>&1
ls -la gpurun_out
