"""GPU parity of the fe_operator_t residual / tangent path against oracle/nonlinear_oracle.py, and a Newton
solve driven through the mirror of fe_operator_t (set_evaluation_point / compute_residual / compute_tangent)."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR, FUNC_SHARP_WAVE
from oracle.nonlinear_oracle import residual_and_tangent

pytestmark = pytest.mark.gpu
WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def _op(fb2, N, p, beta, func, prm):
    s = fb2.FeSpace(N, order=p)
    op = fb2.FeOperator()
    op.create(fb2.csr_format, [False], [False], [0], s, fb2.NonlinearDiscreteIntegration(beta, func, prm))
    return s, op


@pytest.mark.parametrize("N,p,beta,func,prm", [((4, 3, 3), 2, 0.0, FUNC_LINEAR, None), ((4, 3, 3), 2, 0.7, FUNC_SHARP_WAVE, WAVE),
                                               ((6, 5, 4), 1, 1.3, FUNC_LINEAR, None), ((3, 2, 2), 3, 0.4, FUNC_LINEAR, None)])
def test_residual_and_tangent_match_oracle(fb2, N, p, beta, func, prm):
    s, op = _op(fb2, N, p, beta, func, prm)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, func, prm, route=1)                 # fills the oracle's Dirichlet values
    nq = (p + 1) ** 3
    fq = np.random.default_rng(5).standard_normal((s.num_local_cells, nq))
    s.set_source_values(fq)
    u = np.random.default_rng(2).standard_normal(s.num_free_dofs)
    x, y = s.create_dof_values(), s.create_dof_values()
    x.set_entries(u)
    op.apply(x, y)                                               # y = residual(x)
    assert op.get_state() == "residual_computed"
    r_ref, T_ref = residual_and_tangent(o, N, u, beta=beta, f_q=fq)
    assert np.max(np.abs(y.get_entries() - r_ref)) <= 1e-12 * np.max(np.abs(r_ref))
    A = op.get_tangent()
    assert op.get_state() == "assembler_computed"
    irp, ja = A.get_pattern()
    assert np.array_equal(irp - 1, T_ref.indptr) and np.array_equal(ja - 1, T_ref.indices)
    v = A.get_values()
    assert np.max(np.abs(v - T_ref.data)) <= 1e-12 * np.max(np.abs(T_ref.data))
    # a new evaluation point invalidates both
    op.set_evaluation_point(x)
    assert op.get_state() == "created"
    s.set_source_values(None)


def test_newton_on_the_quasilinear_problem(fb2):
    """-div((1 + beta u^2) grad u) = f with u = x + y + z (in the Q2 space): f = -6 beta u.  Newton through the fe_operator
    mirror, GMRES on the non-symmetric tangent; quadratic convergence and the exact discrete solution."""
    N, p, beta = (6, 6, 6), 2, 0.5
    s, op = _op(fb2, N, p, beta, FUNC_LINEAR, None)
    xq = s.get_quadrature_points()
    s.set_source_values(-6.0 * beta * xq.sum(axis=2))
    u, r, du = s.create_dof_values(), s.create_dof_values(), s.create_dof_values()
    u.init(0.0)
    ils = fb2.IterativeLinearSolver(); ils.set_type_from_string("RGMRES")
    ils.set_parameters_from_pl({fb2.ILS_RTOL: 1e-12, fb2.ILS_MAX_NUM_ITERATIONS: 2000, fb2.ILS_MAX_DIM_KRYLOV_BASIS: 60,
                                fb2.ILS_OUTPUT_FREQUENCY: 0})
    norms = []
    for it in range(12):
        op.apply(u, r)
        norms.append(r.nrm2())
        if norms[-1] <= 1e-11 * norms[0]:
            break
        A = op.get_tangent()
        M = fb2.JacobiPreconditioner(); M.create(A); M.numerical_setup()
        ils.set_operators(A, M)
        ils.apply(r, du)
        assert ils.converged()
        u.axpby(-1.0, du, 1.0)                                    # u <- u - T^-1 r
    assert norms[-1] <= 1e-11 * norms[0], norms
    assert len(norms) <= 9
    tail = [n for n in norms if n > 0]
    assert tail[-1] <= 1e-3 * tail[-2] or tail[-1] < 1e-12        # the last step is (super)quadratic
    e = s.error_norms_all(u, FUNC_LINEAR, None)
    assert all(abs(v) < 1e-8 for v in e.values()), e
    s.set_source_values(None)
