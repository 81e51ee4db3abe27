# launch list of the set-up kernels at config 2 (first 160 launches of the bench command)
set -x
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file gpurun_out/c22_launches_setup.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e-assembly > gpurun_out/c22_ncu.log 2>&1
tail -2 gpurun_out/c22_ncu.log
