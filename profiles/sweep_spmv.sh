#!/bin/bash
# sweep of the SpMV tile geometry (rows per tile R, pipeline stages) at Q2 128^3; prints achieved GB/s
for R in 8 16 32 64; do for S in 2 3 4 6 8; do
  out=$(FB2_SPMV_R=$R FB2_SPMV_STAGES=$S python bench.py --cells ${CELLS:-128} --steps 2 --warmup 1 --no-cpu-baseline 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.0f GB/s spmv_ms=%.3f'%(d['value'],d['spmv_ms']))" 2>&1)
  echo "R=$R STAGES=$S $out"
done; done
