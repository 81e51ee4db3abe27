"""CPU tests: pin the oracle against the reference's only numeric golden data for this path
(Tutorials/tutorial_01_poisson_sharp_circular_wave/convergence_plot/uniform_mesh_error_order_{1,2,4},
copied as tests/golden/tutorial_01_uniform_mesh_error.json by tests/golden/make_golden.py) and against
the par_test_poisson acceptance criterion."""
import json
import os

import numpy as np
import pytest

from oracle.oracle import OracleSpace, KIND_POISSON, KIND_VECTOR_LAPLACIAN, FUNC_LINEAR, FUNC_SHARP_WAVE

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "tutorial_01_uniform_mesh_error.json")))
PRM2D = [200.0, 0.7, -0.05, -0.05, 0.0]


@pytest.mark.parametrize("order,row", [(o, r) for o in ("1", "2", "4") for r in range(len(GOLD[o]))])
def test_tutorial_01_golden_energy_error(order, row):
    sqrt_ndofs, err = GOLD[order][row]
    p = int(order)
    N = int(round((sqrt_ndofs + 1) / p))
    s = OracleSpace([N, N], order=p, dims=2)
    assert s.num_free == int(round(sqrt_ndofs)) ** 2          # NUM_DOFS printed by the tutorial
    s.assemble(KIND_POISSON, FUNC_SHARP_WAVE, PRM2D, route=0)
    s.jacobi_setup()
    r = s.cg(prec=1, rtol=1e-14, maxit=50000)
    assert r["converged"]
    h1, _ = s.errors(r["x"], FUNC_SHARP_WAVE, PRM2D)
    assert abs(h1 - err) <= 2e-9 * err                        # the file holds 10 significant digits


@pytest.mark.parametrize("N,p", [((6, 6, 6), 1), ((4, 4, 4), 2), ((3, 3, 3), 3)])
@pytest.mark.parametrize("prec", [0, 1])
def test_par_test_poisson_acceptance_serial(N, p, prec):
    """u = x+y+z, f = 0, CG rtol 1e-12: error norms < 1e-4 (par_test_poisson_driver.f90:417-419,576-586)."""
    s = OracleSpace(N, order=p)
    s.assemble(KIND_POISSON, FUNC_LINEAR, route=0)
    s.jacobi_setup()
    r = s.cg(prec=prec, rtol=1e-12, maxit=5000)
    assert r["converged"]
    h1, l2 = s.errors(r["x"], FUNC_LINEAR)
    assert h1 < 1e-4 and l2 < 1e-4


def test_sizes_of_survey_table():
    s = OracleSpace([10, 10, 10], order=1)
    s.build_pattern(False)
    assert (s.num_free, s.nnz) == (729, 15625)
    s.build_pattern(True)
    assert s.nnz == 8177


def test_coo_route_equals_direct_route_bitwise():
    for sym in (False, True):
        a = OracleSpace([4, 3, 5], order=2); a.assemble(KIND_POISSON, FUNC_SHARP_WAVE, [200, .7, -.05, -.05, -.05], sym=sym, route=0)
        b = OracleSpace([4, 3, 5], order=2); b.assemble(KIND_POISSON, FUNC_SHARP_WAVE, [200, .7, -.05, -.05, -.05], sym=sym, route=1)
        assert np.array_equal(a.irp(), b.irp()) and np.array_equal(a.ja(), b.ja())
        assert np.array_equal(a.val(), b.val()) and np.array_equal(a.rhs(), b.rhs())


def test_closed_form_q1_element_and_row_sums():
    s = OracleSpace([4, 4, 4], order=1)
    elmat, _ = s.element(0)
    h = 0.25
    assert abs(elmat[0, 0] - h / 3.0) < 1e-14 and abs(elmat[0, 7] + h / 12.0) < 1e-14
    assert np.allclose(elmat.sum(axis=1), 0.0, atol=1e-14) and np.allclose(elmat, elmat.T, atol=1e-15)
    s.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
    y = s.spmv(np.ones(s.num_free))
    irp = s.irp()
    interior = np.where(np.diff(irp) == 27)[0]
    # rows whose whole stencil is free sum to zero (constants are in the kernel of the Laplacian)
    g = s.lst_dofs_gids()
    free_cells = np.all(g > 0, axis=1)
    assert interior.size > 0 and free_cells.sum() == 8
    center = [r for r in interior if abs(y[r]) < 1e-13]
    assert len(center) >= 1


def test_vector_laplacian_keeps_explicit_zeros():
    s = OracleSpace([3, 3, 3], order=1, ncomp=3)
    s.assemble(KIND_VECTOR_LAPLACIAN, FUNC_LINEAR, route=0)
    assert s.num_free == 3 * 8 and s.nnz == 24 * 24
    assert np.count_nonzero(s.val() == 0.0) >= 24 * 16


def test_oracle_cg_stopping_criteria_are_consistent():
    """The ten criteria of cg.f90:279-407 on one system: every run converges, the residual-type ones coincide where their
    right-hand sides coincide (x0 = 0 => r(0) = b), and the AND-combinations never stop before their first criterion."""
    from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_SHARP_WAVE
    s = OracleSpace([5, 5, 5], order=2)
    s.assemble(KIND_POISSON, FUNC_SHARP_WAVE, [200.0, 0.7, -0.05, -0.05, -0.05], route=1)
    s.jacobi_setup()
    it = [s.cg(prec=1, rtol=1e-8, maxit=2000, stop=k) for k in range(10)]
    assert all(r["converged"] for r in it)
    n = [r["iters"] for r in it]
    assert n[0] == n[1] and n[2] == n[3] and n[6] == n[7] and n[8] == n[9]
    assert n[6] >= n[4] and n[8] >= n[5]
