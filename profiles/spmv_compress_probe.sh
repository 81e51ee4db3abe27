#!/bin/bash
# uint16 dictionary-compressed column indices (x gathered once per distinct column of a tile) vs plain int32 ja in the
# streaming SpMV; Q2, 1 GPU.  Extra arguments: stage capacities to try with compression.
python - <<'P'
import fempar_b200 as fb2
fb2.init(0)
s = fb2.FeSpace([96] * 3, order=2)
op = fb2.FeAffineOperator()
op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, fb2.func_linear, None, None, fb2.scatter_atomic))
op.compute()
x, y = s.create_dof_values(), s.create_dof_values()
x.init(1.0)
op.get_matrix().apply(x, y)
for ph in ("spmv_R", "spmv_tiles", "spmv_ctas_per_sm", "spmv_compress", "spmv_dict_entries", "spmv_dict_cap"):
    print(ph, op.last_phase_ms(ph))
print("nnz", op.get_matrix().nnz)
P
for cells in ${CELLS:-128 256}; do
  for cfg in "0 2304" "1 2304" "1 2048" "1 1792"; do
    set -- $cfg
    out=$(FB2_SPMV_COMPRESS=$1 FB2_SPMV_CAP=$2 FB2_SPMV_R=32 python bench.py --cells $cells --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.0f GB/s spmv_ms=%.3f pcg_ms/iter=%.3f e2e=%.0f'%(d['value'],d['spmv_ms'],d['pcg_ms_per_iteration'],d['e2e']['value']))" 2>&1)
    echo "Q2 ${cells}^3 compress=$1 cap=$2: $out"
  done
done
