"""GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle on identical meshes.

Bars (BASELINE.json north_star): DOF numbering, irp, ja bit-exact; matrix entries and RHS within
1e-12 relative; CG iteration count within +-1 and same convergence flag.
"""
import numpy as np
import pytest

from oracle.oracle import (OracleSpace, KIND_POISSON, KIND_VECTOR_LAPLACIAN, KIND_ELASTICITY, FUNC_LINEAR,
                           FUNC_SHARP_WAVE)

pytestmark = pytest.mark.gpu

WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]
VAL_RTOL = 1e-12


def _close(a, b, rtol=VAL_RTOL):
    scale = max(np.max(np.abs(b)), 1e-300) if b.size else 1.0
    return np.max(np.abs(a - b)) <= rtol * scale if b.size else True


CASES = [((4, 4, 4), 1, 1), ((10, 10, 10), 1, 1), ((5, 4, 3), 2, 1), ((6, 6, 6), 2, 1), ((3, 4, 2), 3, 1),
         ((1, 1, 1), 2, 1), ((2, 1, 3), 1, 1), ((4, 3, 3), 2, 3), ((3, 3, 3), 1, 3), ((2, 2, 2), 4, 1)]


@pytest.mark.parametrize("N,p,ncomp", CASES)
def test_dof_numbering_bit_exact(fb2, N, p, ncomp):
    s = fb2.FeSpace(N, order=p, num_components=ncomp)
    o = OracleSpace(N, order=p, ncomp=ncomp)
    assert s.num_free_dofs == o.num_free
    assert s.num_dirichlet_dofs == o.num_dirichlet
    assert np.array_equal(s.get_lst_dofs_gids(), o.lst_dofs_gids())


@pytest.mark.parametrize("N,p,ncomp", CASES)
@pytest.mark.parametrize("sym", [False, True])
def test_csr_graph_bit_exact(fb2, N, p, ncomp, sym):
    s = fb2.FeSpace(N, order=p, num_components=ncomp)
    A = fb2.SparseMatrix(s, symmetric_storage=sym)
    o = OracleSpace(N, order=p, ncomp=ncomp)
    o.build_pattern(sym=sym)
    irp, ja = A.get_pattern()
    assert A.nnz == o.nnz
    assert np.array_equal(irp - 1, o.irp())
    assert np.array_equal(ja, o.ja())


def test_csr_graph_matches_reference_faithful_coo_route(fb2):
    """route 0 of the oracle = COO append -> stable sort -> compress, as the reference does it."""
    N, p = (5, 5, 5), 2
    s = fb2.FeSpace(N, order=p)
    o = OracleSpace(N, order=p)
    for sym in (False, True):
        A = fb2.SparseMatrix(s, symmetric_storage=sym)
        o.assemble(KIND_POISSON, FUNC_LINEAR, sym=sym, route=0)
        irp, ja = A.get_pattern()
        assert np.array_equal(irp - 1, o.irp()) and np.array_equal(ja, o.ja())


def _assemble(fb2, N, p, ncomp, kind, func, prm, mat, scatter, sym=False):
    s = fb2.FeSpace(N, order=p, num_components=ncomp)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [sym], [True], [1], s, fb2.DiscreteIntegration(kind, func, prm, mat, scatter))
    op.compute()
    return s, op


@pytest.fixture(params=[0, 1], ids=["asm_fast", "asm_generic"])
def asm_variant(request, fb2):
    """0 = register-blocked contraction + scatter map (where it applies), 1 = generic kernel searching ja."""
    fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(request.param))
    yield request.param
    fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(0))


@pytest.mark.parametrize("N,p", [((4, 4, 4), 1), ((10, 10, 10), 1), ((1, 1, 1), 1), ((2, 1, 3), 2), ((5, 4, 3), 2), ((6, 6, 6), 2),
                                 ((9, 8, 7), 2), ((3, 3, 2), 3)])
@pytest.mark.parametrize("func,prm", [(FUNC_LINEAR, None), (FUNC_SHARP_WAVE, WAVE)])
@pytest.mark.parametrize("scatter", [0, 1])
def test_poisson_assembly_matches_oracle(fb2, asm_variant, N, p, func, prm, scatter):
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, func, prm, None, scatter)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, func, prm, route=1)
    assert _close(s.get_fixed_dof_values(), o.fixed())
    assert _close(op.get_matrix().get_values(), o.val())
    assert _close(op.get_translation().get_entries(), o.rhs())


def test_poisson_assembly_symmetric_storage(fb2):
    N, p = (5, 5, 5), 2
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0, sym=True)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, sym=True, route=0)
    assert _close(op.get_matrix().get_values(), o.val())
    assert _close(op.get_translation().get_entries(), o.rhs())
    # symmetric-storage SpMV (csr_sparse_matrix.f90:597-623)
    x = np.random.default_rng(0).standard_normal(s.num_free_dofs)
    vx, vy = s.create_dof_values(), s.create_dof_values()
    vx.set_entries(x)
    op.get_matrix().apply(vx, vy)
    assert _close(vy.get_entries(), o.spmv(x), 1e-12)


@pytest.mark.parametrize("kind,okind,mat", [(1, KIND_VECTOR_LAPLACIAN, None), (2, KIND_ELASTICITY, [5.0 / 18.0, 5.0 / 12.0])])
@pytest.mark.parametrize("scatter", [0, 1])
def test_vector_assembly_matches_oracle(fb2, asm_variant, kind, okind, mat, scatter):
    N, p = (4, 3, 3), 2
    s, op = _assemble(fb2, N, p, 3, kind, FUNC_LINEAR, None, mat, scatter)
    o = OracleSpace(N, order=p, ncomp=3)
    o.assemble(okind, FUNC_LINEAR, mat=mat, route=1)
    assert _close(op.get_matrix().get_values(), o.val())
    assert _close(op.get_translation().get_entries(), o.rhs())


@pytest.mark.parametrize("N,p", [((10, 10, 10), 1), ((6, 6, 6), 2), ((3, 3, 3), 3)])
def test_spmv_and_blas1_match_oracle(fb2, N, p):
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    n = s.num_free_dofs
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    vx, vy, vz = s.create_dof_values(), s.create_dof_values(), s.create_dof_values()
    vx.set_entries(x)
    op.get_matrix().apply(vx, vz)
    ref = o.spmv(x)
    assert np.max(np.abs(vz.get_entries() - ref)) <= 1e-13 * np.max(np.abs(ref))
    vy.set_entries(y)
    assert abs(vx.dot(vy) - float(x @ y)) <= 1e-12 * np.sqrt(float(x @ x) * float(y @ y))
    assert abs(vx.nrm2() - np.linalg.norm(x)) <= 1e-13 * np.linalg.norm(x)
    vy.axpby(2.5, vx, -0.5)
    assert np.allclose(vy.get_entries(), 2.5 * x - 0.5 * y, rtol=1e-15, atol=1e-15)
    vy.axpby(3.0, vx, 0.0)
    assert np.array_equal(vy.get_entries(), 3.0 * x)
    vz.entrywise_product(vx, vy)
    assert np.array_equal(vz.get_entries(), x * (3.0 * x))
    vz.scal(-2.0, vx)
    assert np.array_equal(vz.get_entries(), -2.0 * x)
    vz.copy(vx)
    vz.entrywise_invert()
    assert np.array_equal(vz.get_entries(), 1.0 / x)
    vz.init(7.0)
    assert np.all(vz.get_entries() == 7.0)
    d = s.create_dof_values()
    op.get_matrix().extract_diagonal(d)
    dinv = o.jacobi_setup()
    assert np.allclose(1.0 / d.get_entries(), dinv, rtol=1e-12)


@pytest.mark.parametrize("N,p,func,prm,prec", [((10, 10, 10), 1, FUNC_SHARP_WAVE, WAVE, "jacobi"),
                                                ((10, 10, 10), 1, FUNC_LINEAR, None, "identity"),
                                                ((8, 8, 8), 2, FUNC_LINEAR, None, "jacobi"),
                                                ((6, 6, 6), 2, FUNC_SHARP_WAVE, WAVE, "jacobi"),
                                                ((4, 4, 4), 3, FUNC_LINEAR, None, "jacobi")])
@pytest.mark.parametrize("rtol", [1e-6, 1e-12])
def test_pcg_matches_oracle(fb2, N, p, func, prm, prec, rtol):
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, func, prm, None, 0)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, func, prm, route=1)
    M = None
    if prec == "jacobi":
        M = fb2.JacobiPreconditioner(); M.create(op); M.symbolic_setup(); M.numerical_setup()
        o.jacobi_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_type_from_string("CG")
    ils.set_parameters_from_pl({fb2.ILS_RTOL: rtol, fb2.ILS_MAX_NUM_ITERATIONS: 5000, fb2.ILS_TRACK_CONVERGENCE_HISTORY: True})
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    r = o.cg(prec=1 if M else 0, rtol=rtol, maxit=5000, history=True)
    assert ils.converged() == r["converged"]
    assert abs(ils.get_num_iterations() - r["iters"]) <= 1
    k = min(ils.get_num_iterations(), r["iters"]) // 2
    assert np.allclose(ils.history[:k], r["history"][:k], rtol=1e-6)
    xs = x.get_entries()
    assert np.max(np.abs(xs - r["x"])) <= 100 * rtol * max(1.0, np.max(np.abs(r["x"])))
    # the reference's own acceptance criterion (par_test_poisson_driver.f90:576-586): error norms < 1e-4 for u = x+y+z
    if func == FUNC_LINEAR and rtol == 1e-12:
        h1, l2 = s.error_norms(x, func, prm)
        assert h1 < 1e-4 and l2 < 1e-4
        oh1, ol2 = o.errors(r["x"], func, prm)
        assert oh1 < 1e-4 and ol2 < 1e-4


def test_pcg_host_buffers_and_max_iterations(fb2):
    N, p = (6, 6, 6), 2
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    o.jacobi_setup()
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-10, fb2.ILS_MAX_NUM_ITERATIONS: 2000})
    ils.set_operators(op.get_tangent(), M)
    b = op.get_translation().get_entries()
    x = np.zeros_like(b)
    ils.apply(b, x)                         # host buffers through fb2_pcg_solve_host
    r = o.cg(prec=1, rtol=1e-10, maxit=2000)
    assert ils.converged() and abs(ils.get_num_iterations() - r["iters"]) <= 1
    assert np.max(np.abs(x - r["x"])) < 1e-8
    # max-iteration exit leaves did_converge = .false. (cg.f90:214)
    ils.set_parameters_from_pl({fb2.ILS_MAX_NUM_ITERATIONS: 5})
    ils.apply(b, x)
    assert not ils.converged() and ils.get_num_iterations() == 5
    r5 = o.cg(prec=1, rtol=1e-10, maxit=5)
    assert np.allclose(x, r5["x"], rtol=1e-9, atol=1e-12)
    # RES_RHS criterium
    ils.set_parameters_from_pl({fb2.ILS_MAX_NUM_ITERATIONS: 2000, fb2.ILS_STOPPING_CRITERIUM: fb2.res_rhs})
    ils.apply(b, x)
    r2 = o.cg(prec=1, rtol=1e-10, maxit=2000, stop=1)
    assert abs(ils.get_num_iterations() - r2["iters"]) <= 1


def test_tutorial_01_config1_end_to_end(fb2):
    """config 1: tutorial_01 3-D Q1 10^3 with the sharp-wave data; energy error equals the oracle's."""
    N, p = (10, 10, 10), 1
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0)
    assert s.num_free_dofs == 729 and op.get_matrix().nnz == 15625
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-13, fb2.ILS_MAX_NUM_ITERATIONS: 5000})
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=0)
    o.jacobi_setup()
    r = o.cg(prec=1, rtol=1e-13, maxit=5000)
    h1, l2 = s.error_norms(x, FUNC_SHARP_WAVE, WAVE)
    oh1, ol2 = o.errors(r["x"], FUNC_SHARP_WAVE, WAVE)
    assert abs(h1 - oh1) <= 1e-9 * oh1 and abs(l2 - ol2) <= 1e-9 * ol2


STOPS = ["RES_RES", "RES_RHS", "RES_NRMGIVEN_RHS_NRMGIVEN", "RES_NRMGIVEN_RES_NRMGIVEN", "DELTA_RHS", "DELTA_DELTA",
         "DELTA_RHS_AND_RES_RES", "DELTA_RHS_AND_RES_RHS", "DELTA_DELTA_AND_RES_RES", "DELTA_DELTA_AND_RES_RHS"]


@pytest.mark.parametrize("stop", range(10), ids=STOPS)
@pytest.mark.parametrize("prec", ["jacobi", "identity"])
def test_pcg_all_ten_stopping_criteria(fb2, stop, prec):
    """cg_supports_stopping_criteria (cg.f90:240-255): every criterion stops at the oracle's iteration (+-1) with the
    oracle's error estimate; init_convergence_data / update_convergence_data (cg.f90:279-407)."""
    N, p = (6, 6, 6), 2
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    M = None
    if prec == "jacobi":
        M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
        o.jacobi_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-8, fb2.ILS_MAX_NUM_ITERATIONS: 2000,
                                fb2.ILS_STOPPING_CRITERIUM: STOPS[stop], fb2.ILS_TRACK_CONVERGENCE_HISTORY: True})
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    r = o.cg(prec=1 if M else 0, rtol=1e-8, maxit=2000, stop=stop, history=True)
    assert ils.converged() and r["converged"]
    assert abs(ils.get_num_iterations() - r["iters"]) <= 1
    k = min(ils.get_num_iterations(), r["iters"]) // 2
    assert np.allclose(ils.history[:k], r["history"][:k], rtol=1e-6)
    assert np.max(np.abs(x.get_entries() - r["x"])) <= 1e-6 * np.max(np.abs(r["x"]))


def test_fe_affine_operator_state_machine_and_residual(fb2):
    """fe_affine_operator_t: compute is a no-op once assembler_computed, force_compute re-assembles into the same pattern
    (sbm_fe_affine_operator.i90:15-57); apply(x,y) = residual A x - b at the evaluation point (sbm_fe_operator.i90:435-467)."""
    N, p = (5, 4, 3), 2
    s = fb2.FeSpace(N, order=p)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0))
    assert op.get_state() == "created" and op.is_linear()
    op.compute()
    assert op.get_state() == "assembler_computed"
    v0 = op.get_matrix().get_values().copy()
    op.get_matrix().init(7.0)                      # tamper: compute() must NOT re-assemble, force_compute() must
    op.compute()
    assert np.all(op.get_matrix().get_values() == 7.0)
    op.force_compute()
    assert np.allclose(op.get_matrix().get_values(), v0, rtol=1e-14, atol=1e-300)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    x = np.random.default_rng(5).standard_normal(s.num_free_dofs)
    vx, vy = s.create_dof_values(), s.create_dof_values()
    vx.set_entries(x)
    op.apply(vx, vy)
    ref = o.spmv(x) - o.rhs()
    assert np.max(np.abs(vy.get_entries() - ref)) <= 1e-12 * np.max(np.abs(ref))
    op.apply_add(vx, vy)
    assert np.max(np.abs(vy.get_entries() - 2 * ref)) <= 1e-12 * np.max(np.abs(ref))


def test_determinism_of_reductions_and_colored_scatter(fb2):
    """the fused <Ap,p> and the vector reductions sum per-block partials in a fixed order, so two solves on the
    same matrix are bitwise identical; the colored scatter variant (plain read-modify-write on one colour class at a time)
    makes the assembled values bitwise reproducible too (the atomic variant is only reproducible to rounding)."""
    N, p = (12, 11, 10), 2
    s = fb2.FeSpace(N, order=p)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, fb2.scatter_colored))
    op.compute()
    v1, b1 = op.get_matrix().get_values().copy(), op.get_translation().get_entries().copy()
    op.force_compute()
    assert np.array_equal(op.get_matrix().get_values(), v1) and np.array_equal(op.get_translation().get_entries(), b1)
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    sols, hists = [], []
    for _ in range(2):
        ils = fb2.IterativeLinearSolver()
        ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-10, fb2.ILS_MAX_NUM_ITERATIONS: 2000,
                                    fb2.ILS_TRACK_CONVERGENCE_HISTORY: True})
        ils.set_operators(op.get_tangent(), M)
        x = s.create_dof_values()
        ils.apply(op.get_translation(), x)
        sols.append(x.get_entries().copy()); hists.append(ils.history.copy())
    assert np.array_equal(sols[0], sols[1]) and np.array_equal(hists[0], hists[1])


@pytest.mark.parametrize("N,p", [((9, 8, 7), 2), ((12, 12, 12), 1), ((5, 5, 4), 3)])
def test_spmv_kernel_variants_agree(fb2, N, p, monkeypatch):
    """variant 0 = TMA-streamed tiles, 1 = warp per row, 2 = streamed tiles with dictionary-compressed uint16 column indices:
    2 forms the same products in the same order (bit-identical y and fused <Ap,p>, same PCG history), 1 agrees to rounding.
    (natural row map for both: the transposed map of variant 0 keeps y bit for bit but sums the fused <Ap,p> partials in
    another order -- tests/test_gpu_midsize.py::test_midsize_spmv_variants_bitwise covers it)"""
    monkeypatch.setenv("FB2_SPMV_TRANSPOSE", "0")
    ys, hist = {}, {}
    lib = fb2._lib.lib()
    try:
        for v in (0, 2, 1):
            fb2._lib.check(lib.fb2_set_spmv_variant(v))
            s, op = _assemble(fb2, N, p, 1, KIND_POISSON, FUNC_SHARP_WAVE, WAVE, None, 1)   # coloured scatter: the same matrix bits every time
            x = np.random.default_rng(11).standard_normal(s.num_free_dofs)
            vx, vy = s.create_dof_values(), s.create_dof_values()
            vx.set_entries(x)
            op.get_matrix().apply(vx, vy)
            ys[v] = vy.get_entries()
            M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
            ils = fb2.IterativeLinearSolver()
            ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-10, fb2.ILS_MAX_NUM_ITERATIONS: 500,
                                        fb2.ILS_TRACK_CONVERGENCE_HISTORY: True})
            ils.set_operators(op.get_tangent(), M)
            sol = s.create_dof_values()
            ils.apply(op.get_translation(), sol)
            assert ils.converged()
            hist[v] = (ils.get_num_iterations(), sol.get_entries())
            if v == 2:
                assert op.last_phase_ms("spmv_dict_cap") > 0          # the compressed stream was really built
            op.free()
    finally:
        fb2._lib.check(lib.fb2_set_spmv_variant(0))
    assert np.array_equal(ys[0], ys[2])
    assert hist[0][0] == hist[2][0] and np.array_equal(hist[0][1], hist[2][1])
    assert _close(ys[1], ys[0]) and abs(hist[1][0] - hist[0][0]) <= 1


def _wave_source(x, alpha=200.0, r=0.7, c=-0.05):
    """f = -lap(u) of the tutorial_01 sharp circular wave, evaluated on the host (what a user integrand would do)."""
    d2 = (x - c) ** 2
    t1 = d2.sum(-1); s = np.sqrt(t1); t2 = s - r; w = alpha ** 2 * t2 ** 2 + 1.0
    res = 0.0
    for k in range(3):
        res = res - (alpha / (s * w) - alpha * d2[..., k] / (t1 * s * w) - 2 * alpha ** 3 * d2[..., k] * t2 / (t1 * w * w))
    return res


@pytest.mark.parametrize("N,p", [((6, 5, 4), 1), ((5, 4, 3), 2), ((3, 3, 2), 3)])
def test_user_source_term_sampled_at_quadrature_points(fb2, asm_variant, N, p):
    """integrand seam for an arbitrary f: quadrature points from the device, f evaluated on the host, samples handed back;
    with the tutorial's f this must reproduce the built-in analytic source, and the RHS is linear in the samples."""
    s, op = _assemble(fb2, N, p, 1, KIND_POISSON, FUNC_SHARP_WAVE, WAVE, None, 0)
    b_ref = op.get_translation().get_entries()
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=0)
    assert _close(b_ref, o.rhs())
    xq = s.get_quadrature_points()
    assert xq.shape == (s.num_local_cells, (p + 1) ** 3, 3) and xq.min() > 0.0 and xq.max() < 1.0
    f = _wave_source(xq)
    s.set_source_values(f)
    op.force_compute()
    b1 = op.get_translation().get_entries()
    assert np.max(np.abs(b1 - b_ref)) <= 1e-11 * np.max(np.abs(b_ref))
    s.set_source_values(np.zeros_like(f))
    op.force_compute()
    b0 = op.get_translation().get_entries()                      # Dirichlet lifting only
    s.set_source_values(-2.5 * f)
    op.force_compute()
    b2 = op.get_translation().get_entries()
    assert np.max(np.abs((b2 - b0) + 2.5 * (b1 - b0))) <= 1e-11 * np.max(np.abs(b_ref))
    s.set_source_values(None)                                    # back to the analytic selector
    op.force_compute()
    assert _close(op.get_translation().get_entries(), b_ref, 1e-13)
    with pytest.raises(fb2.Fb2Error):
        s.set_source_values(np.zeros(3))
    op.free()


@pytest.mark.parametrize("N,p,func,prm", [((6, 6, 6), 2, FUNC_SHARP_WAVE, WAVE), ((5, 4, 3), 3, FUNC_SHARP_WAVE, WAVE), ((8, 8, 8), 1, FUNC_LINEAR, None)])
@pytest.mark.parametrize("exponent", [None, 3])
def test_eleven_error_norms_match_oracle(fb2, N, p, func, prm, exponent):
    """error_norms_scalar_t%compute for the norm types par_test_poisson checks (par_test_poisson_driver.f90:557-589)"""
    s, op = _assemble(fb2, N, p, 1, fb2.poisson, func, prm, None, 0)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, func, prm, route=1)
    o.jacobi_setup()
    r = o.cg(prec=1, rtol=1e-12, maxit=5000)
    x = s.create_dof_values()
    x.set_entries(r["x"])
    got = s.error_norms_all(x, func, prm, exponent)
    want, _ = o.errors_all(r["x"], func, prm, exponent)
    assert list(got) == list(want) == list(fb2.ERROR_NORM_NAMES)
    scale = max(abs(v) for v in want.values())
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-9 * max(abs(want[k]), 1e-12 * scale) + 1e-13, (k, got[k], want[k])
    h1, l2 = s.error_norms(x, func, prm)
    assert abs(h1 - want["h1_seminorm"]) <= 1e-9 * abs(want["h1_seminorm"]) + 1e-13 and abs(l2 - want["l2_norm"]) <= 1e-9 * abs(want["l2_norm"]) + 1e-13
    if func == FUNC_LINEAR:                       # the reference's acceptance bar
        assert all(abs(v) < 1e-4 for v in got.values())


def test_convergence_log_in_reference_format(fb2):
    """header / per-iteration line / footer of base_iterative_linear_solver.f90:584-676, ILS_OUTPUT_FREQUENCY and ILS_LUOUT"""
    import io
    import re
    s, op = _assemble(fb2, (6, 6, 6), 2, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None, 0)
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    o = OracleSpace((6, 6, 6), order=2)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    o.jacobi_setup()
    es = r"  [ -]\d\.\d{9}E[+-]\d{2}"
    for stop, ostop, two in ((fb2.res_res, 0, False), (fb2.delta_rhs_and_res_rhs, 7, True)):
        r = o.cg(prec=1, rtol=1e-8, maxit=500, stop=ostop, history=True)
        buf = io.StringIO()
        ils = fb2.IterativeLinearSolver(); ils.set_type_from_string("CG")
        ils.set_parameters_from_pl({fb2.ILS_RTOL: 1e-8, fb2.ILS_MAX_NUM_ITERATIONS: 500, fb2.ILS_STOPPING_CRITERIUM: stop,
                                    fb2.ILS_OUTPUT_FREQUENCY: 5, fb2.ILS_LUOUT: buf})
        ils.set_operators(op.get_tangent(), M)
        x = s.create_dof_values()
        ils.apply(op.get_translation(), x)
        lines = buf.getvalue().splitlines()
        it = ils.get_num_iterations()
        assert abs(it - r["iters"]) <= 1
        cols = 4 if two else 2
        assert lines[0] == "CG:   " + "Iteration" + "".join("  " + "%15s" % t for t in ("Error Estimate", "Tolerance") * (cols // 2))
        body = lines[1:-2]
        want_iters = [i for i in range(1, it + 1) if i % 5 == 0 or i == it]
        assert [int(ln[6:15]) for ln in body] == want_iters
        for ln in body:
            assert re.fullmatch(r"CG:   [ \d]{9}" + es * cols, ln), ln
        k = want_iters[0]
        assert abs(float(body[0][15:33]) - r["history"][k - 1]) <= 1e-6 * r["history"][k - 1]
        assert re.fullmatch(r"CG converged to " + es + (r" [ -]\d\.\d{9}E[+-]\d{2}" if two else "") + r"  in  [ \d]{9}  iterations\. ", lines[-2]), lines[-2]
        assert lines[-1].startswith("Last iteration error estimate" + ("s: " if two else ": "))
    ils.set_parameters_from_pl({fb2.ILS_OUTPUT_FREQUENCY: 0})
    buf.truncate(0); buf.seek(0)
    ils.apply(op.get_translation(), x)
    assert buf.getvalue() == ""
