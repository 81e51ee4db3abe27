"""Full-size (BASELINE.json configs 2, 4, 5) property tests of the CUDA path, sized for one B200 (180 GB).

The oracle cannot run these sizes in seconds, so the checks are size-independent properties:
  * closed-form sizes: NUM_DOFS = (pN-1)^3 per component, nnz of the full graph = (sum over lattice nodes of the
    per-direction coupling widths)^3, both independent of any numbering;
  * consistency of the discretisation with u = x+y+z (in the FE space): A u_I = b to rounding, i.e. integration, Dirichlet
    elimination and scatter of every one of the 16.8 M cells are right (a wrong slot or a lost atomic shows up here);
  * symmetry x.Ay = y.Ax of the assembled operator through the SpMV kernel;
  * symmetric-storage SpMV == full-storage SpMV (two independent graphs, two independent kernels);
  * the reference's own acceptance test (par_test_poisson_driver.f90:576-586): PCG to rtol 1e-12, error norms < 1e-4;
  * re-assembly (force_compute, sbm_fe_operator.i90:568-639) reproduces the same values (idempotence).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _nnz_full(N, p, ncomp=1):
    """Couplings of the free nodes of a p-th order lattice on N cells per direction: per direction, a node at lattice
    position m (1..pN-1 free) couples with the free nodes of the 1 or 2 cells that contain it."""
    def width(m):
        lo = (m // p) * p if m % p else m - p          # first node of the left-most cell containing m
        hi = (m // p + 1) * p if m % p else m + p
        lo, hi = max(lo, 1), min(hi, p * N - 1)
        return hi - lo + 1
    w = sum(width(m) for m in range(1, p * N))
    return ncomp * ncomp * w ** 3


def _free_gb():
    import torch
    return torch.cuda.mem_get_info()[0] / 2 ** 30


@pytest.mark.parametrize("N,p,ncomp,kind,need_gb", [(256, 2, 1, 0, 125), (128, 3, 1, 0, 105), (128, 2, 3, 1, 150), (128, 2, 3, 2, 150)],
                         ids=["config2_Q2_256", "config5_Q3_128", "config4_vecQ2_128", "config4_elasticity_128"])
def test_fullsize_assembly_properties(fb2, N, p, ncomp, kind, need_gb):
    if _free_gb() < need_gb:
        pytest.skip("needs %d GB of free HBM" % need_gb)
    s = fb2.FeSpace([N] * 3, order=p, num_components=ncomp)
    assert s.num_free_dofs == ncomp * (p * N - 1) ** 3
    op = fb2.FeAffineOperator()
    mat = [5.0 / 18.0, 5.0 / 12.0] if kind == 2 else None
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(kind, fb2.func_linear, None, mat, 0))
    A = op.get_matrix()
    assert A.nnz == _nnz_full(N, p, ncomp)
    op.compute()
    b = op.get_translation()
    u, Au, y, Ay = (s.create_dof_values() for _ in range(4))
    s.interpolate(fb2.func_linear, None, u)
    A.apply(u, Au)
    nb = b.nrm2()
    Au.axpby(-1.0, b, 1.0)
    # 1e-12 relative to the size of the terms that cancel (|A_ij u_j| ~ |diag| * |u|)
    assert Au.nrm2() <= 1e-12 * max(nb, 1.0) * 10, (Au.nrm2(), nb)
    # symmetry through the SpMV kernel
    rng = np.random.default_rng(0)
    n = s.num_free_dofs
    y.set_entries(rng.standard_normal(n))
    u.set_entries(rng.standard_normal(n))
    A.apply(u, Au)
    A.apply(y, Ay)
    a1, a2 = y.dot(Au), u.dot(Ay)
    assert abs(a1 - a2) <= 1e-10 * max(abs(a1), Au.nrm2() * y.nrm2() * 1e-3)
    # idempotent re-assembly: compute() zeroes and refills the same pattern
    op.force_compute()
    A.apply(u, Ay)
    Ay.axpby(-1.0, Au, 1.0)
    assert Ay.nrm2() <= 1e-13 * Au.nrm2()
    for v in (u, Au, y, Ay):
        v.free()
    op.free()
    s.free()


def test_fullsize_sym_storage_spmv_equals_full(fb2):
    N, p = 160, 2
    if _free_gb() < 50:
        pytest.skip("needs 50 GB of free HBM")
    s = fb2.FeSpace([N] * 3, order=p)
    ops = []
    for sym in (False, True):
        op = fb2.FeAffineOperator()
        op.create(fb2.csr_format, [sym], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, fb2.func_linear))
        op.compute()
        ops.append(op)
    # upper triangle incl. diagonal
    assert ops[1].get_matrix().nnz == (ops[0].get_matrix().nnz + s.num_free_dofs) // 2
    x, y0, y1 = (s.create_dof_values() for _ in range(3))
    x.set_entries(np.random.default_rng(1).standard_normal(s.num_free_dofs))
    ops[0].get_matrix().apply(x, y0)
    ops[1].get_matrix().apply(x, y1)
    y1.axpby(-1.0, y0, 1.0)
    assert y1.nrm2() <= 1e-12 * y0.nrm2()
    for v in (x, y0, y1):
        v.free()
    for op in ops:
        op.free()
    s.free()


def test_fullsize_acceptance_q2_128_pcg(fb2):
    """par_test_poisson's criterion at 128^3 Q2 (16.6 M DOFs): Jacobi-PCG to rtol 1e-12 converges and reproduces x+y+z."""
    N, p = 128, 2
    s = fb2.FeSpace([N] * 3, order=p)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(fb2.poisson, fb2.func_linear))
    op.compute()
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    ils = fb2.IterativeLinearSolver()
    ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-12, fb2.ILS_MAX_NUM_ITERATIONS: 5000,
                                fb2.ILS_OUTPUT_FREQUENCY: 100})
    ils.check_every = 32
    ils.set_operators(op.get_tangent(), M)
    x = s.create_dof_values()
    ils.apply(op.get_translation(), x)
    assert ils.converged() and ils.get_num_iterations() < 5000
    h1, l2 = s.error_norms(x, fb2.func_linear, None)
    assert h1 < 1e-4 and l2 < 1e-4
    norms = s.error_norms_all(x, fb2.func_linear, None)          # the eleven norms of par_test_poisson_driver.f90:557-589
    assert all(abs(v) < 1e-4 for v in norms.values()), norms
    u = s.create_dof_values()
    s.interpolate(fb2.func_linear, None, u)
    u.axpby(-1.0, x, 1.0)
    assert u.nrm2() <= 1e-8 * x.nrm2()
    for v in (x, u):
        v.free()
    op.free()
    s.free()
