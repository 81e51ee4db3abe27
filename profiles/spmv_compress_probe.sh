#!/bin/bash
# uint16 dictionary-compressed column indices (x gathered once per distinct column of a tile) vs plain int32 ja in the
# streaming SpMV; Q2, 1 GPU.
for cells in ${CELLS:-128 256}; do
  for v in 0 2; do
    out=$(python bench.py --cells $cells --steps 3 --warmup 3 --no-cpu-baseline --spmv-variant $v 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.0f GB/s spmv_ms=%.3f pcg_ms/iter=%.3f e2e=%.0f'%(d['value'],d['spmv_ms'],d['pcg_ms_per_iteration'],d['e2e']['value']))" 2>&1)
    echo "Q2 ${cells}^3 spmv-variant=$v: $out"
  done
done
