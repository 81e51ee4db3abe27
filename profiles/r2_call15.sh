set -x
for tb in 512 128; do
FB2_SPMV_TPR_BASE=$tb timeout 300 python -m pytest tests/test_gpu_midsize.py -q -x -k "spmv or midsize_serial" > gpurun_out/c15_pytest_tpr$tb.log 2>&1; echo "rc=$?" >> gpurun_out/c15_pytest_tpr$tb.log
FB2_SPMV_TPR_BASE=$tb timeout 600 python profiles/spmv_sweep.py 256 2 > gpurun_out/c15_spmv_sweep_tpr${tb}_q2_256.txt 2>&1
FB2_SPMV_TPR_BASE=$tb timeout 600 python profiles/spmv_sweep.py 128 3 > gpurun_out/c15_spmv_sweep_tpr${tb}_q3_128.txt 2>&1
done
