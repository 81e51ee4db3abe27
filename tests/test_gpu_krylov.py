"""GPU parity of the other Krylov solvers sharing the CG primitives: ICG, RICHARDSON, RGMRES,
FGMRES, LGMRES against the numpy restatement in oracle/krylov_oracle.py (iteration count +-1, convergence history,
solution)."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_SHARP_WAVE  # noqa: F401
from oracle import krylov_oracle as K

pytestmark = pytest.mark.gpu
WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]
STOPS = ["RES_RES", "RES_RHS", "RES_NRMGIVEN_RHS_NRMGIVEN", "RES_NRMGIVEN_RES_NRMGIVEN", "DELTA_RHS", "DELTA_DELTA",
         "DELTA_RHS_AND_RES_RES", "DELTA_RHS_AND_RES_RHS", "DELTA_DELTA_AND_RES_RES", "DELTA_DELTA_AND_RES_RHS"]


@pytest.fixture(scope="module")
def system(fb2):
    N, p = (6, 6, 6), 2
    s = fb2.FeSpace(N, order=p)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0))
    op.compute()
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    dinv = o.jacobi_setup().copy()
    A = sp.csr_matrix((o.val().copy(), (o.ja() - 1).copy(), o.irp().copy()), shape=(o.num_free, o.num_free))  # o is freed on return
    return dict(s=s, op=op, M=M, A=lambda x: A @ x, Mj=lambda r: dinv * r, Mi=lambda r: r.copy(), b=o.rhs().copy())


def _run(fb2, sysd, name, prec, pl):
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl(dict({fb2.ILS_TYPE: name, fb2.ILS_TRACK_CONVERGENCE_HISTORY: True}, **pl))
    ils.set_operators(sysd["op"].get_tangent(), sysd["M"] if prec else None)
    x = sysd["s"].create_dof_values()
    ils.apply(sysd["op"].get_translation(), x)
    return ils, x.get_entries()


def _compare(ils, x, r, xtol=1e-6):
    assert ils.converged() == r["converged"]
    assert abs(ils.get_num_iterations() - r["iters"]) <= 1
    k = min(ils.get_num_iterations(), r["iters"]) * 2 // 3
    assert np.allclose(ils.history[:k], r["history"][:k], rtol=1e-5)
    assert np.max(np.abs(x - r["x"])) <= xtol * np.max(np.abs(r["x"]))


@pytest.mark.parametrize("stop", range(10), ids=STOPS)
def test_icg_all_stopping_criteria(fb2, system, stop):
    ils, x = _run(fb2, system, "ICG", True, {fb2.ILS_RTOL: 1e-8, fb2.ILS_MAX_NUM_ITERATIONS: 2000, fb2.ILS_STOPPING_CRITERIUM: STOPS[stop]})
    _compare(ils, x, K.icg(system["A"], system["Mj"], system["b"], rtol=1e-8, maxit=2000, stop=stop))


@pytest.mark.parametrize("stop", [0, 1])
def test_richardson(fb2, system, stop):
    pl = {fb2.ILS_RTOL: 1e-3, fb2.ILS_MAX_NUM_ITERATIONS: 400, fb2.ILS_STOPPING_CRITERIUM: STOPS[stop], fb2.ILS_RELAXATION: 0.9}
    ils, x = _run(fb2, system, "RICHARDSON", True, pl)
    r = K.richardson(system["A"], system["Mj"], system["b"], rtol=1e-3, maxit=400, stop=stop, relaxation=0.9)
    _compare(ils, x, r, xtol=1e-9)
    # max-iteration exit
    ils, x = _run(fb2, system, "RICHARDSON", True, dict(pl, **{fb2.ILS_MAX_NUM_ITERATIONS: 7}))
    r = K.richardson(system["A"], system["Mj"], system["b"], rtol=1e-3, maxit=7, stop=stop, relaxation=0.9)
    assert ils.get_num_iterations() == 7 and not ils.converged() and np.allclose(x, r["x"], rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("name,fn", [("RGMRES", K.rgmres), ("FGMRES", K.fgmres), ("LGMRES", K.lgmres)])
@pytest.mark.parametrize("strat", ["ICGSRO", "MGSRO"])
@pytest.mark.parametrize("dkry", [8, 30])
@pytest.mark.parametrize("stop", [2, 3])
@pytest.mark.parametrize("prec", [True, False])
def test_gmres_family(fb2, system, name, fn, strat, dkry, stop, prec):
    pl = {fb2.ILS_RTOL: 1e-8, fb2.ILS_MAX_NUM_ITERATIONS: 500, fb2.ILS_STOPPING_CRITERIUM: STOPS[stop],
          fb2.ILS_MAX_DIM_KRYLOV_BASIS: dkry, fb2.ILS_ORTHONORM_STRATEGY: strat}
    ils, x = _run(fb2, system, name, prec, pl)
    r = fn(system["A"], system["Mj"] if prec else system["Mi"], system["b"], rtol=1e-8, maxit=500, stop=stop, dkrymax=dkry,
           strat=K.ICGSRO if strat == "ICGSRO" else K.MGSRO)
    _compare(ils, x, r)


@pytest.mark.parametrize("stop", [0, 1])
def test_lgmres_true_residual_criteria(fb2, system, stop):
    pl = {fb2.ILS_RTOL: 1e-8, fb2.ILS_MAX_NUM_ITERATIONS: 500, fb2.ILS_STOPPING_CRITERIUM: STOPS[stop], fb2.ILS_MAX_DIM_KRYLOV_BASIS: 12}
    ils, x = _run(fb2, system, "LGMRES", True, pl)
    _compare(ils, x, K.lgmres(system["A"], system["Mj"], system["b"], rtol=1e-8, maxit=500, stop=stop, dkrymax=12))


def test_default_stopping_criteria_and_unsupported(fb2, system):
    ils = fb2.IterativeLinearSolver()
    ils.set_type_from_string("RGMRES")
    assert ils.params[fb2.ILS_STOPPING_CRITERIUM] == "RES_NRMGIVEN_RES_NRMGIVEN"
    with pytest.raises(fb2.Fb2Error):
        ils.set_type_from_string("PARDISO")                     # not an iterative solver key
    ils, _ = None, None
    with pytest.raises(fb2.Fb2Error):                           # rgmres_supports_stopping_criteria (rgmres.f90:358-366)
        _run(fb2, system, "RGMRES", True, {fb2.ILS_STOPPING_CRITERIUM: "RES_RES"})


@pytest.mark.parametrize("stop", [0, 1])
@pytest.mark.parametrize("strat", ["ICGSRO", "MGSRO"])
@pytest.mark.parametrize("dkry", [60, 10])
def test_lfom(fb2, system, stop, strat, dkry):
    """dkry = 10 exercises the restart exactly as the reference writes it (g(1) stays the initial norm, lfom.f90:247): both
    sides stagnate the same way and stop at max_num_iterations."""
    pl = {fb2.ILS_RTOL: 1e-9, fb2.ILS_MAX_NUM_ITERATIONS: 60, fb2.ILS_STOPPING_CRITERIUM: STOPS[stop],
          fb2.ILS_MAX_DIM_KRYLOV_BASIS: dkry, fb2.ILS_ORTHONORM_STRATEGY: strat}
    ils, x = _run(fb2, system, "LFOM", True, pl)
    r = K.lfom(system["A"], system["Mj"], system["b"], rtol=1e-9, maxit=60, stop=stop, dkrymax=dkry,
               strat=K.ICGSRO if strat == "ICGSRO" else K.MGSRO)
    assert ils.converged() == r["converged"] and abs(ils.get_num_iterations() - r["iters"]) <= 1
    k = min(ils.get_num_iterations(), r["iters"]) * 2 // 3
    assert np.allclose(ils.history[:k], r["history"][:k], rtol=1e-5)
    if r["converged"]:
        assert np.max(np.abs(x - r["x"])) <= 1e-6 * np.max(np.abs(r["x"]))


@pytest.mark.parametrize("prec", [True, False])
def test_minres_spd(fb2, system, prec):
    pl = {fb2.ILS_RTOL: 1e-10, fb2.ILS_MAX_NUM_ITERATIONS: 500}
    ils, x = _run(fb2, system, "MINRES", prec, pl)
    r = K.minres(system["A"], system["Mj"] if prec else system["Mi"], system["b"], rtol=1e-10, maxit=500)
    _compare(ils, x, r)
    assert int(system["op"].last_phase_ms("minres_istop")) == r["istop"]
    assert abs(system["op"].last_phase_ms("minres_Anorm") - r["Anorm"]) <= 1e-8 * r["Anorm"]
    # iteration limit: istop = 5, did_converge = .false.
    ils, x = _run(fb2, system, "MINRES", prec, dict(pl, **{fb2.ILS_MAX_NUM_ITERATIONS: 6}))
    r = K.minres(system["A"], system["Mj"] if prec else system["Mi"], system["b"], rtol=1e-10, maxit=6)
    assert ils.get_num_iterations() == 6 and not ils.converged() and r["istop"] == 5
    assert np.allclose(x, r["x"], rtol=1e-10, atol=1e-13)


def test_minres_symmetric_indefinite_host_matrix(fb2):
    """MINRES is the reference's solver for symmetric INDEFINITE operators: A - sigma I with sigma inside the spectrum,
    handed over as a host-assembled CSR matrix."""
    o = OracleSpace((5, 5, 5), order=2)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    n = o.num_free
    A = sp.csr_matrix((o.val().copy(), (o.ja() - 1).copy(), o.irp().copy()), shape=(n, n))
    ev = np.linalg.eigvalsh(A.toarray())
    k = n // 4 + int(np.argmax(np.diff(ev[n // 4:3 * n // 4])))     # widest spectral gap in the middle half
    sigma = 0.5 * (ev[k] + ev[k + 1])
    As = (A - sigma * sp.identity(n)).tocsr()
    As.sort_indices()
    b = o.rhs().copy()
    Ad = fb2.SparseMatrix(host_csr=(As.indptr + 1, As.indices + 1, As.data))
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "MINRES", fb2.ILS_RTOL: 1e-10, fb2.ILS_MAX_NUM_ITERATIONS: 3000,
                                fb2.ILS_TRACK_CONVERGENCE_HISTORY: True})
    ils.set_operators(Ad, None)
    vb, vx = fb2.ScalarArray(n=n), fb2.ScalarArray(n=n)
    vb.set_entries(b)
    ils.apply(vb, vx)
    r = K.minres(lambda v: As @ v, lambda v: v.copy(), b, rtol=1e-10, maxit=3000)
    assert ils.converged() and r["converged"] and abs(ils.get_num_iterations() - r["iters"]) <= 2
    xs = np.linalg.solve(As.toarray(), b)
    assert np.max(np.abs(vx.get_entries() - xs)) <= 1e-6 * np.max(np.abs(xs))


def _np_pcg(A, dinv, b, x0, rtol, maxit):
    """cg.f90:101-238 with a non-zero initial solution (RES_RES: ||r_i|| <= rtol ||r_0||), numpy on the oracle's CSR."""
    x = x0.copy()
    r = b - A @ x
    z = dinv * r
    p = z.copy()
    rz = r @ z
    tol = rtol * np.sqrt(r @ r)
    it = 0
    while True:
        it += 1
        Ap = A @ p
        alpha = rz / (Ap @ p)
        x += alpha * p
        r -= alpha * Ap
        if np.sqrt(r @ r) <= tol or it >= maxit:
            return x, it
        z = dinv * r
        beta = 1.0 / rz
        rz = r @ z
        beta *= rz
        p = z + beta * p


@pytest.mark.parametrize("ils_type", ["CG", "RGMRES", "MINRES", "RICHARDSON"])
def test_set_initial_solution(fb2, ils_type):
    """base_iterative_linear_solver_t%set_initial_solution: the solve starts from a copy of the given vector (default zero)."""
    import scipy.sparse as sp
    from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_SHARP_WAVE
    N, p = (6, 5, 4), 2
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, [200.0, 0.7, -0.05, -0.05, -0.05], route=0)
    A = sp.csr_matrix((o.val().copy(), o.ja().copy() - 1, o.irp().copy()), shape=(o.num_free, o.num_free))
    dinv, b = o.jacobi_setup().copy(), o.rhs().copy()
    s = fb2.FeSpace(N, order=p)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, FUNC_SHARP_WAVE,
                                                                                [200.0, 0.7, -0.05, -0.05, -0.05], None, 1))
    op.compute()
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    x0 = np.random.default_rng(4).standard_normal(o.num_free)
    v0 = s.create_dof_values(); v0.set_entries(x0)
    ils = fb2.IterativeLinearSolver()
    relax = {fb2.ILS_RELAXATION: 0.9} if ils_type == "RICHARDSON" else {}
    ils.set_parameters_from_pl(dict({fb2.ILS_TYPE: ils_type, fb2.ILS_RTOL: 1e-9, fb2.ILS_MAX_NUM_ITERATIONS: 3000}, **relax))
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    it_zero, sol_zero = ils.get_num_iterations(), x.get_entries()
    ils.set_initial_solution(v0)
    v0.init(123.0)                                               # the solver keeps its own copy
    ils.apply(op.get_translation(), x)
    assert ils.converged()
    # every solver converges to the same solution from the new start; CG's RES_RES is relative to the NEW initial residual
    # (the other types stop on their own estimates of a preconditioned norm: checked through the solution)
    assert np.linalg.norm(x.get_entries() - sol_zero) <= 1e-5 * np.linalg.norm(sol_zero)
    if ils_type == "CG":
        assert np.linalg.norm(b - A @ x.get_entries()) <= 1.01e-9 * np.linalg.norm(b - A @ x0)
    if ils_type == "CG":
        xr, itr = _np_pcg(A, dinv, b, x0, 1e-9, 3000)
        assert abs(ils.get_num_iterations() - itr) <= 1 and np.max(np.abs(x.get_entries() - xr)) < 1e-8
        xh = np.zeros(o.num_free)
        ils.apply(b, xh)                                         # host-buffer entry uses it too
        assert np.max(np.abs(xh - xr)) < 1e-8
    # starting from the converged solution: nothing left to do beyond the first iteration
    v0.set_entries(sol_zero); ils.set_initial_solution(v0)
    ils.set_parameters_from_pl({fb2.ILS_RTOL: 0.0, fb2.ILS_ATOL: 1e-5 * max(np.linalg.norm(b), np.linalg.norm(dinv * b))})
    ils.apply(op.get_translation(), x)
    assert ils.converged() and ils.get_num_iterations() <= 1
    ils.set_initial_solution(None)                               # back to zero
    ils.set_parameters_from_pl({fb2.ILS_RTOL: 1e-9, fb2.ILS_ATOL: 0.0})
    ils.apply(op.get_translation(), x)
    assert ils.get_num_iterations() == it_zero and np.array_equal(x.get_entries(), sol_zero)
    op.free()
