#!/bin/bash
# chunked-CTA SpMV: tile rows R, stage capacity (entries), stages, tiles per CTA; Q2 ${CELLS:-128}^3, 1 GPU
for cfg in "32 2304 2 32" "32 2304 2 64" "16 1152 4 64" "16 1152 3 64" "16 1536 3 64" "32 3072 2 32" "32 2048 2 32" "64 4608 2 16" "16 1152 2 64" "32 1792 3 32" "8 768 6 128"; do
  set -- $cfg
  out=$(FB2_SPMV_R=$1 FB2_SPMV_CAP=$2 FB2_SPMV_STAGES=$3 FB2_SPMV_CHUNK=$4 python bench.py --cells ${CELLS:-128} --steps 2 --warmup 2 --no-cpu-baseline 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.0f GB/s spmv_ms=%.3f'%(d['value'],d['spmv_ms']))" 2>&1)
  echo "R=$1 cap=$2 stages=$3 chunk=$4: $out"
done
