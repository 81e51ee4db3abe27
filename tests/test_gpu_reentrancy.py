"""the C-ABI is re-entrant across handles.  Solver work vectors, the device-resident convergence state, the
initial solution and the host-buffer staging vectors live in the matrix handle, so operators of different size can be used
alternately (a block system, a time stepper with mass + stiffness) without re-allocation or cross-talk."""
import numpy as np
import pytest

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR, FUNC_SHARP_WAVE

pytestmark = pytest.mark.gpu
WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def _system(fb2, N, p, func, prm):
    s = fb2.FeSpace(N, order=p)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, func, prm))
    op.compute()
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, func, prm, route=1)
    o.jacobi_setup()
    return s, op, M, o


def _solver(fb2, op, M, rtol=1e-10, track=True, stop=None):
    ils = fb2.IterativeLinearSolver(); ils.set_type_from_string("CG")
    ils.set_parameters_from_pl({fb2.ILS_RTOL: rtol, fb2.ILS_MAX_NUM_ITERATIONS: 500, fb2.ILS_TRACK_CONVERGENCE_HISTORY: track})
    if stop is not None:
        ils.set_parameters_from_pl({fb2.ILS_STOPPING_CRITERIUM: stop})
    ils.set_operators(op.get_tangent(), M)
    return ils


def test_two_operators_interleaved(fb2):
    s1, op1, M1, o1 = _system(fb2, (10, 9, 8), 1, FUNC_SHARP_WAVE, WAVE)
    s2, op2, M2, o2 = _system(fb2, (6, 6, 6), 2, FUNC_LINEAR, None)
    assert s1.num_free_dofs != s2.num_free_dofs
    r1 = o1.cg(prec=1, rtol=1e-10, maxit=500, history=True)
    r2 = o2.cg(prec=1, rtol=1e-10, maxit=500, history=True)
    ils1, ils2, ils1b = _solver(fb2, op1, M1), _solver(fb2, op2, M2), _solver(fb2, op1, M1)
    x1, x2 = s1.create_dof_values(), s2.create_dof_values()
    for rnd in range(3):                                   # A1, A2, A1 (second solver object), A2 through host buffers, ...
        ils1.apply(op1.get_translation(), x1)
        ils2.apply(op2.get_translation(), x2)
        assert abs(ils1.get_num_iterations() - r1["iters"]) <= 1 and abs(ils2.get_num_iterations() - r2["iters"]) <= 1
        assert np.max(np.abs(x1.get_entries() - r1["x"])) <= 1e-8 * np.max(np.abs(r1["x"]))
        assert np.max(np.abs(x2.get_entries() - r2["x"])) <= 1e-8 * np.max(np.abs(r2["x"]))
        k = min(len(ils1.history), len(r1["history"]))
        assert np.max(np.abs(ils1.history[:k] - r1["history"][:k]) / r1["history"][:k]) <= 1e-6
        ils1b.apply(op1.get_translation(), x1)
        assert abs(ils1b.get_num_iterations() - r1["iters"]) <= 1
        xh2 = np.zeros(s2.num_free_dofs)
        ils2.apply(op2.get_translation().get_entries(), xh2)        # host buffers (fb2_pcg_solve_host)
        xh1 = np.zeros(s1.num_free_dofs)
        ils1.apply(op1.get_translation().get_entries(), xh1)
        assert np.max(np.abs(xh2 - r2["x"])) <= 1e-8 * np.max(np.abs(r2["x"]))
        assert np.max(np.abs(xh1 - r1["x"])) <= 1e-8 * np.max(np.abs(r1["x"]))


def test_initial_solution_is_per_operator(fb2):
    """an initial solution set for one operator must not leak into the solves of another one of the SAME size"""
    s1, op1, M1, o1 = _system(fb2, (6, 6, 6), 2, FUNC_SHARP_WAVE, WAVE)
    s2, op2, M2, o2 = _system(fb2, (6, 6, 6), 2, FUNC_LINEAR, None)
    # RES_RHS: the tolerance is relative to |b|, so a warm start from the converged solution stops at once
    r1 = o1.cg(prec=1, rtol=1e-10, maxit=500, stop=1)
    r2 = o2.cg(prec=1, rtol=1e-10, maxit=500, stop=1)
    ils1, ils2 = _solver(fb2, op1, M1, track=False, stop=fb2.res_rhs), _solver(fb2, op2, M2, track=False, stop=fb2.res_rhs)
    x1, x2 = s1.create_dof_values(), s2.create_dof_values()
    ils1.apply(op1.get_translation(), x1)
    n_cold = ils1.get_num_iterations()
    ils1.set_initial_solution(x1)                          # warm start of operator 1: converged at once
    ils1.apply(op1.get_translation(), x1)
    assert ils1.get_num_iterations() <= 2 < n_cold
    ils2.apply(op2.get_translation(), x2)                  # operator 2 still starts from zero
    assert abs(ils2.get_num_iterations() - r2["iters"]) <= 1
    assert np.max(np.abs(x2.get_entries() - r2["x"])) <= 1e-8 * np.max(np.abs(r2["x"]))
    ils1.set_initial_solution(None)
    ils1.apply(op1.get_translation(), x1)
    assert abs(ils1.get_num_iterations() - r1["iters"]) <= 1
