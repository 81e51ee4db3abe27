# allocation-free dot: the new test alone in a fresh process, then the whole GPU suite (the new file is collected last)
set -x
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c26_pytest_no_alloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/c26_pytest_no_alloc_alone.log
tail -4 gpurun_out/c26_pytest_no_alloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c26_pytest_all.log 2>&1; echo "rc=$?" >> gpurun_out/c26_pytest_all.log
tail -4 gpurun_out/c26_pytest_all.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c26_smoke.log 2>&1; tail -2 gpurun_out/c26_smoke.log
