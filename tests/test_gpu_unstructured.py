"""GPU tests of host-numbered FE spaces on unstructured / distorted conforming hex meshes: the reference's
lst_dofs_gids and vertex coordinates are uploaded, the device builds the graph, integrates with per-cell Jacobians, solves."""
import numpy as np
import pytest

from oracle import hex_oracle as ho
from oracle.oracle import OracleSpace, KIND_POISSON, KIND_ELASTICITY, FUNC_LINEAR, FUNC_SHARP_WAVE

pytestmark = pytest.mark.gpu
WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def warp(X):
    x, y, z = X[..., 0], X[..., 1], X[..., 2]
    return np.stack([x + 0.06 * np.sin(np.pi * x) * np.sin(2 * np.pi * y) * (1 + z),
                     y + 0.05 * np.sin(np.pi * y) * np.sin(np.pi * z) * (1 - 0.5 * x),
                     z + 0.04 * np.sin(np.pi * z) * np.cos(np.pi * x) * y], axis=-1)


def _close(a, b, rtol=1e-12):
    return np.max(np.abs(a - b)) <= rtol * max(np.max(np.abs(b)), 1e-300)


@pytest.fixture(params=[0, 1], ids=["asm_fast", "asm_generic"], autouse=True)
def asm_variant(request, fb2):
    """0 = sum-factorised kernel with one explicit scatter table per cell, 1 = generic kernel searching ja."""
    fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(request.param))
    yield request.param
    fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(0))


def _operator(fb2, s, sym, kind=KIND_POISSON, func=FUNC_LINEAR, prm=None, mat=None):
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [sym], [True], [1], s, fb2.DiscreteIntegration(kind, func, prm, mat, fb2.scatter_atomic))
    op.compute()
    return op


@pytest.mark.parametrize("p,sym", [(1, False), (2, False), (2, True), (3, False)])
def test_uploaded_structured_space_equals_device_numbered_space(fb2, p, sym):
    """the same mesh through both doors: graph bit-exact, values / RHS / error norms to 1e-12."""
    N = 5
    s0 = fb2.FeSpace([N] * 3, order=p)
    s0.interpolate_dirichlet_values(FUNC_SHARP_WAVE, WAVE)
    op0 = _operator(fb2, s0, sym, func=FUNC_SHARP_WAVE, prm=WAVE)
    gids = s0.get_lst_dofs_gids()
    s1 = fb2.HostNumberedFeSpace(p, 1, gids, ho.structured_cell_coords(N), s0.num_free_dofs, s0.num_dirichlet_dofs)
    assert (s1.num_cells, s1.num_local_cells, s1.num_free_dofs, s1.num_dirichlet_dofs) == (N ** 3, N ** 3, s0.num_free_dofs,
                                                                                           s0.num_dirichlet_dofs)
    s1.interpolate_dirichlet_values(FUNC_SHARP_WAVE, WAVE)
    assert _close(s1.get_fixed_dof_values(), s0.get_fixed_dof_values(), 1e-14)
    op1 = _operator(fb2, s1, sym, func=FUNC_SHARP_WAVE, prm=WAVE)
    i0, j0 = op0.get_matrix().get_pattern()
    i1, j1 = op1.get_matrix().get_pattern()
    assert np.array_equal(i0, i1) and np.array_equal(j0, j1)
    assert _close(op1.get_matrix().get_values(), op0.get_matrix().get_values())
    assert _close(op1.get_translation().get_entries(), op0.get_translation().get_entries(), 1e-11)
    x0, x1 = s0.create_dof_values(), s1.create_dof_values()
    s0.interpolate(FUNC_SHARP_WAVE, WAVE, x0); s1.interpolate(FUNC_SHARP_WAVE, WAVE, x1)
    assert _close(x1.get_entries(), x0.get_entries(), 1e-14)
    e0, e1 = s0.error_norms(x0, FUNC_SHARP_WAVE, WAVE), s1.error_norms(x1, FUNC_SHARP_WAVE, WAVE)
    assert abs(e0[0] - e1[0]) <= 1e-12 * e0[0] and abs(e0[1] - e1[1]) <= 1e-12 * e0[1]
    for o in (op0, op1):
        o.free()


@pytest.mark.parametrize("p,sym", [(1, False), (2, False), (2, True)])
def test_distorted_mesh_matches_hex_oracle_and_patch_test(fb2, p, sym):
    """non-affine cells, cells handed over in a shuffled order: values against oracle/hex_oracle.py, then the isoparametric
    patch test (u = x+y+z reproduced exactly) through Jacobi-PCG and the device error norms."""
    N = 4
    o = OracleSpace([N] * 3, order=p)
    perm = np.random.default_rng(5).permutation(N ** 3)
    gids = o.lst_dofs_gids().copy()[perm]
    X = ho.structured_cell_coords(N, warp)[perm]
    s = fb2.HostNumberedFeSpace(p, 1, gids, X, o.num_free, o.num_dirichlet)
    s.interpolate_dirichlet_values(FUNC_LINEAR)
    xn = ho.node_coordinates(p, X).sum(-1)
    fixed = np.zeros(o.num_dirichlet)
    fixed[-gids[gids < 0] - 1] = xn[gids < 0]
    assert _close(s.get_fixed_dof_values(), fixed, 1e-14)
    op = _operator(fb2, s, sym)
    A, rhs = ho.assemble(p, 1, gids, X, o.num_free, fixed, sym=sym)
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, A.indptr) and np.array_equal(ja - 1, A.indices)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * np.abs(rhs).max()
    if sym:
        op.free()
        return
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-13, fb2.ILS_MAX_NUM_ITERATIONS: 2000})
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    assert ils.converged()
    exact = np.zeros(o.num_free)
    exact[gids[gids > 0] - 1] = xn[gids > 0]
    assert np.max(np.abs(x.get_entries() - exact)) < 1e-9
    h1, l2 = s.error_norms(x, FUNC_LINEAR)
    assert h1 < 1e-8 and l2 < 1e-9
    rh1, rl2 = ho.error_norms(p, gids, X, x.get_entries(), fixed, lambda q: q.sum(-1), lambda q: np.ones_like(q))
    assert abs(h1 - rh1) < 1e-9 and abs(l2 - rl2) < 1e-9
    op.free()


def test_distorted_elasticity_matches_hex_oracle(fb2):
    N, p, lam, mu = 3, 1, 1.3, 0.7
    o = OracleSpace([N] * 3, order=p, ncomp=3)
    gids = o.lst_dofs_gids().copy()
    X = ho.structured_cell_coords(N, warp)
    s = fb2.HostNumberedFeSpace(p, 3, gids, X, o.num_free, o.num_dirichlet)
    op = _operator(fb2, s, False, kind=KIND_ELASTICITY, mat=[lam, mu])     # compute() interpolates u = x+y+z per component
    xn = np.tile(ho.node_coordinates(p, X).sum(-1), (1, 3))
    fixed = np.zeros(o.num_dirichlet)
    fixed[-gids[gids < 0] - 1] = xn[gids < 0]
    assert _close(s.get_fixed_dof_values(), fixed, 1e-14)
    A, rhs = ho.assemble(p, 3, gids, X, o.num_free, fixed, lam=lam, mu=mu)
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, A.indptr) and np.array_equal(ja - 1, A.indices)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * np.abs(rhs).max()
    op.free()


def test_local_cells_only_and_ghost_dofs_are_skipped(fb2):
    """par layout of a host-numbered space: trailing ghost cells are not integrated, ids above num_free_dofs are skipped."""
    N, p = 3, 1
    o = OracleSpace([N] * 3, order=p)
    gids = o.lst_dofs_gids().copy()
    X = ho.structured_cell_coords(N, warp)
    nloc = 20
    used = np.unique(gids[:nloc][gids[:nloc] > 0])
    remap = np.zeros(o.num_free + 1, dtype=np.int64)
    remap[used] = np.arange(1, used.size + 1)
    ghost_only = np.setdiff1d(np.arange(1, o.num_free + 1), used)
    remap[ghost_only] = used.size + np.arange(1, ghost_only.size + 1)
    g2 = np.where(gids > 0, remap[np.clip(gids, 0, None)], gids).astype(np.int32)
    s = fb2.HostNumberedFeSpace(p, 1, g2, X, used.size, o.num_dirichlet, num_local_cells=nloc)
    s.interpolate_dirichlet_values(FUNC_LINEAR)
    op = _operator(fb2, s, False)
    A, rhs = ho.assemble(p, 1, g2[:nloc], X[:nloc], used.size, s.get_fixed_dof_values())
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, A.indptr) and np.array_equal(ja - 1, A.indices)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * max(np.abs(rhs).max(), 1.0)
    op.free()


def test_bad_arguments_fail_loudly(fb2):
    g = np.ones((2, 8), dtype=np.int32)
    with pytest.raises(fb2.Fb2Error):
        fb2.HostNumberedFeSpace(1, 1, g, np.zeros((2, 8, 2)), 1, 0)             # coordinates not [cells, 8, 3]
    with pytest.raises(fb2.Fb2Error):
        fb2.HostNumberedFeSpace(2, 1, g, np.zeros((2, 8, 3)), 1, 0)             # 8 ids per cell is not Q2
    g[0, 0] = -3
    with pytest.raises(fb2.Fb2Error):
        fb2.HostNumberedFeSpace(1, 1, g, np.zeros((2, 8, 3)), 1, 2)             # Dirichlet id beyond the declared count


@pytest.mark.parametrize("p", [1, 2, 3])
def test_y_junction_unstructured_topology(fb2, p):
    """a mesh no lattice describes (three blocks around a valence-3 line): graph + values against the hex oracle, patch test
    through Jacobi-PCG, interpolation and error norms on the device."""
    X = ho.y_junction_mesh(3, 2)
    gids, nf, ndir = ho.number_by_coordinates(p, X)
    s = fb2.HostNumberedFeSpace(p, 1, gids, X, nf, ndir)
    op = _operator(fb2, s, False)
    xn = ho.node_coordinates(p, X).sum(-1)
    fixed = np.zeros(ndir)
    fixed[-gids[gids < 0] - 1] = xn[gids < 0]
    assert _close(s.get_fixed_dof_values(), fixed, 1e-14)
    A, rhs = ho.assemble(p, 1, gids, X, nf, fixed)
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, A.indptr) and np.array_equal(ja - 1, A.indices)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * np.abs(rhs).max()
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-13, fb2.ILS_MAX_NUM_ITERATIONS: 2000})
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    assert ils.converged()
    exact = np.zeros(nf)
    exact[gids[gids > 0] - 1] = xn[gids > 0]
    assert np.max(np.abs(x.get_entries() - exact)) < 1e-9
    xi = s.create_dof_values()
    s.interpolate(FUNC_LINEAR, None, xi)
    assert _close(xi.get_entries(), exact, 1e-14)
    h1, l2 = s.error_norms(x, FUNC_LINEAR)
    assert h1 < 1e-8 and l2 < 1e-9
    op.free()


def test_y_junction_elasticity(fb2):
    X = ho.y_junction_mesh(2, 2)
    gids, nf, ndir = ho.number_by_coordinates(2, X, ncomp=3)
    s = fb2.HostNumberedFeSpace(2, 3, gids, X, nf, ndir)
    op = _operator(fb2, s, False, kind=KIND_ELASTICITY, mat=[5.0 / 18.0, 5.0 / 12.0])
    A, rhs = ho.assemble(2, 3, gids, X, nf, s.get_fixed_dof_values(), lam=5.0 / 18.0, mu=5.0 / 12.0)
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, A.indptr) and np.array_equal(ja - 1, A.indices)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * np.abs(rhs).max()
    op.free()


@pytest.mark.parametrize("p", [1, 2])
def test_user_source_term_on_distorted_mesh(fb2, p):
    """quadrature points of a host-numbered space follow the trilinear map; a host-evaluated source gives the hex oracle's RHS."""
    X = ho.y_junction_mesh(3, 2)
    gids, nf, ndir = ho.number_by_coordinates(p, X)
    s = fb2.HostNumberedFeSpace(p, 1, gids, X, nf, ndir)
    xq = s.get_quadrature_points()
    assert np.max(np.abs(xq - ho.cell_geometry(p, X)[0])) < 1e-14
    src = lambda x: np.sin(3.0 * x[..., 0]) * np.cos(2.0 * x[..., 1]) + x[..., 2] ** 2
    s.set_source_values(src(xq))
    op = _operator(fb2, s, False)
    A, rhs = ho.assemble(p, 1, gids, X, nf, s.get_fixed_dof_values(), source=src)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * np.abs(rhs).max()
    op.free()


@pytest.mark.parametrize("p", [1, 2])
def test_y_junction_vector_laplacian(fb2, p):
    from oracle.oracle import KIND_VECTOR_LAPLACIAN
    X = ho.y_junction_mesh(2, 2)
    gids, nf, ndir = ho.number_by_coordinates(p, X, ncomp=3)
    s = fb2.HostNumberedFeSpace(p, 3, gids, X, nf, ndir)
    op = _operator(fb2, s, False, kind=KIND_VECTOR_LAPLACIAN)
    A, rhs = ho.assemble(p, 3, gids, X, nf, s.get_fixed_dof_values(), vector_laplacian=True)
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, A.indptr) and np.array_equal(ja - 1, A.indices)
    assert _close(op.get_matrix().get_values(), A.data)
    assert np.max(np.abs(op.get_translation().get_entries() - rhs)) <= 1e-12 * np.abs(rhs).max()
    op.free()
