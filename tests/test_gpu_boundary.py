"""GPU tests of the drop-in seams for host-built FEMPAR objects: a CSR matrix the reference assembled
itself, the element-wise assembler fed in batches, and a dof_import_t supplied by the host."""
import numpy as np
import pytest

from oracle.oracle import OracleSpace, KIND_POISSON, KIND_ELASTICITY, FUNC_LINEAR, FUNC_SHARP_WAVE

pytestmark = pytest.mark.gpu
WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def _close(a, b, rtol=1e-12):
    return np.max(np.abs(a - b)) <= rtol * max(np.max(np.abs(b)), 1e-300)


@pytest.mark.parametrize("sym", [False, True])
def test_host_assembled_csr_spmv_jacobi_pcg(fb2, sym):
    """csr_sparse_matrix_t arrays (irp(1)=1, ja, val) of the reference -> device: apply, extract_diagonal, Jacobi-PCG."""
    N, p = (6, 5, 4), 2
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, sym=sym, route=0)
    A = fb2.SparseMatrix(symmetric_storage=sym, host_csr=(o.irp() + 1, o.ja(), o.val()))
    assert A.nnz == o.nnz and A.num_rows == o.num_free
    irp, ja = A.get_pattern()
    assert np.array_equal(irp - 1, o.irp()) and np.array_equal(ja, o.ja()) and np.array_equal(A.get_values(), o.val())
    n = o.num_free
    x = np.random.default_rng(3).standard_normal(n)
    vx, vy = fb2.ScalarArray(n=n), fb2.ScalarArray(n=n)
    vx.set_entries(x)
    A.apply(vx, vy)
    assert _close(vy.get_entries(), o.spmv(x))
    if not sym:
        M = fb2.JacobiPreconditioner(); M.create(A); M.numerical_setup()
        assert np.allclose(M.inverse_diagonal.get_entries(), o.jacobi_setup(), rtol=1e-14)
        ils = fb2.IterativeLinearSolver()
        ils.set_parameters_from_pl({fb2.ILS_TYPE: "CG", fb2.ILS_RTOL: 1e-10, fb2.ILS_MAX_NUM_ITERATIONS: 1000})
        ils.set_operators(A, M)
        b = o.rhs()
        sol = np.zeros(n)
        ils.apply(b, sol)
        r = o.cg(prec=1, rtol=1e-10, maxit=1000)
        assert ils.converged() and abs(ils.get_num_iterations() - r["iters"]) <= 1
        assert np.max(np.abs(sol - r["x"])) < 1e-8
    A.free()


def test_host_csr_rejects_bad_graph(fb2):
    irp = np.array([1, 3, 5], dtype=np.int32)
    with pytest.raises(fb2.Fb2Error):
        fb2.SparseMatrix(host_csr=(irp, np.array([2, 1, 1, 2], dtype=np.int32), np.ones(4)))      # not ascending
    with pytest.raises(fb2.Fb2Error):
        fb2.SparseMatrix(host_csr=(irp, np.array([1, 2, 1, 3], dtype=np.int32), np.ones(4)))      # column out of range


@pytest.mark.parametrize("N,p,ncomp,kind,mat", [((5, 4, 3), 2, 1, KIND_POISSON, None), ((4, 4, 4), 1, 1, KIND_POISSON, None),
                                                ((3, 3, 2), 2, 3, KIND_ELASTICITY, [5.0 / 18.0, 5.0 / 12.0])])
@pytest.mark.parametrize("device_bc", [True, False])
@pytest.mark.parametrize("sym", [False, True])
def test_element_wise_assembler_batches(fb2, N, p, ncomp, kind, mat, device_bc, sym):
    """assembler_t seam: element matrices computed on the host (here: by the oracle's integrate_galerkin restatement) are
    scattered by the device; equals the oracle's COO->sort->compress assembly.  device_bc = fe_cell_iterator%assembly
    semantics (elimination on the device) vs pure assembler semantics (host already eliminated)."""
    func, prm = (FUNC_SHARP_WAVE, WAVE) if ncomp == 1 else (FUNC_LINEAR, None)
    o = OracleSpace(N, order=p, ncomp=ncomp)
    o.assemble(kind, func, prm, mat=mat, sym=sym, route=0)
    gids = o.lst_dofs_gids()[:o.num_local_cells]
    nd = gids.shape[1]
    fixed = o.fixed()
    s = fb2.FeSpace(N, order=p, num_components=ncomp)
    A = fb2.SparseMatrix(s, symmetric_storage=sym)
    rhs = s.create_dof_values()
    A.init(0.0)
    ncell = gids.shape[0]
    for lo in range(0, ncell, 17):                      # ragged batches
        hi = min(lo + 17, ncell)
        em, ev = [], []
        for c in range(lo, hi):
            E, f = o.element(c, kind, func, prm, mat)
            if not device_bc:                           # sbm_fe_cell_iterator.i90:1610-1637 on the host
                for j in range(nd):
                    if gids[c, j] < 0:
                        f = f - E[:, j] * fixed[-gids[c, j] - 1]
            em.append(np.asfortranarray(E).ravel(order="F")); ev.append(f)
        A.assembly(gids[lo:hi], np.array(em), np.array(ev), rhs, fixed if device_bc else None)
    assert _close(A.get_values(), o.val())
    assert _close(rhs.get_entries(), o.rhs())
    A.assembly(gids[:0].reshape(0, nd), None, None, rhs)  # empty batch is a no-op
    assert _close(rhs.get_entries(), o.rhs())


@pytest.mark.parametrize("nranks,count", [(2, 1), (8, 4), (8, 7), (16, 4)])
def test_peer_allreduce_protocol_on_one_device(fb2, nranks, count):
    """The NVLink peer-memory scalar all-reduce (mailboxes, epoch flags, parity double-buffering) with every rank played by one
    block of a cooperative launch: 200 consecutive all-reduces, every rank must hold the bit-identical rank-ordered sum."""
    import ctypes as C
    err = C.c_double(-1.0)
    fb2._lib.check(fb2._lib.lib().fb2_peer_allreduce_selftest(nranks, count, 200, C.byref(err)))
    assert err.value == 0.0
