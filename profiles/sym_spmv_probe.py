"""SpMV time of symmetric (upper-triangular) storage vs full storage: python profiles/sym_spmv_probe.py [cells]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import fempar_b200 as fb  # noqa: E402
fb.init(0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 128
s = fb.FeSpace([N] * 3, order=2)
for sym in (False, True):
    op = fb.FeAffineOperator()
    op.create(fb.csr_format, [sym], [True], [1], s, fb.DiscreteIntegration(fb.poisson, fb.func_linear))
    op.compute()
    A = op.get_matrix()
    x, y = s.create_dof_values(), s.create_dof_values()
    x.set_entries(np.random.default_rng(0).standard_normal(s.num_free_dofs))
    ts = []
    for _ in range(6):
        A.apply(x, y)
        ts.append(op.last_phase_ms("spmv"))
    t = min(ts[1:])
    n, nnz = s.num_free_dofs, A.nnz
    print("Q2 %d^3 sym=%s nnz=%d: %.3f ms, stored-bytes rate %.0f GB/s, full-matrix-equivalent %.0f GB/s" % (
        N, sym, nnz, t, (nnz * 12 + n * 24) / t / 1e6, ((2 * nnz - n if sym else nnz) * 12 + n * 24) / t / 1e6), flush=True)
    op.free()
