set -x
timeout 300 python -m pytest tests/test_gpu_nonlinear.py -q -x > gpurun_out/c10_pytest_nonlinear.log 2>&1; echo "rc=$?" >> gpurun_out/c10_pytest_nonlinear.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/c10_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c10_pytest_all.log
