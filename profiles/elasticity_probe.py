"""Assembly time of vector Q2 operators (config 4) with the fast and the generic kernel: python profiles/elasticity_probe.py [cells]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fempar_b200 as fb  # noqa: E402
fb.init(0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
s = fb.FeSpace([N] * 3, order=2, num_components=3)
for kind, name, mat in [(fb.vector_laplacian, "vector_laplacian", None), (fb.elasticity, "elasticity", [5.0 / 18.0, 5.0 / 12.0])]:
    op = fb.FeAffineOperator()
    op.create(fb.csr_format, [False], [True], [1], s, fb.DiscreteIntegration(kind, fb.func_linear, None, mat, 0))
    for variant in (0, 1):
        fb._lib.check(fb._lib.lib().fb2_set_assembly_variant(variant))
        ts = []
        for _ in range(4):
            op.force_compute()
            ts.append(op.last_phase_ms("integrate"))
        t = min(ts[1:])
        print("%s Q2 x3 %d^3 (%d DOFs, %d nnz) variant %d: %.2f ms = %.0f MDOF/s" % (name, N, s.num_free_dofs, op.get_matrix().nnz,
                                                                                 variant, t, s.num_free_dofs / t / 1e3), flush=True)
    fb._lib.check(fb._lib.lib().fb2_set_assembly_variant(0))
    op.free()
