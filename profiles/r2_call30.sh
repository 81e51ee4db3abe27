# SpMV: L2 eviction hints on the bulk copies, A/B at config 2 and config 5 sizes
set -x
timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c30_sweep_q2_256.txt 2>&1
cat gpurun_out/c30_sweep_q2_256.txt
timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c30_sweep_q3_128.txt 2>&1
cat gpurun_out/c30_sweep_q3_128.txt
