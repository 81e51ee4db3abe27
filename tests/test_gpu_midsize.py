"""Mid-size GPU parity: sizes at which every code path of the full-size runs is exercised (interior position class of the
scatter/gather maps, ragged 32x32 traversal tiles, row-owner pencils with ragged ends, SpMV pattern classes, > 1 SpMV CTA
chunk) and the oracle still answers in seconds.  Entry-wise bars: every matrix entry above 1e-3 of the
largest is compared RELATIVELY to 1e-12, the small ones absolutely against 1e-12 * max."""
import numpy as np
import pytest

from oracle.oracle import OracleSpace, KIND_POISSON, KIND_VECTOR_LAPLACIAN, KIND_ELASTICITY, FUNC_LINEAR, FUNC_SHARP_WAVE

pytestmark = pytest.mark.gpu
WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def entrywise_close(v, ov, rtol=1e-12):
    assert v.shape == ov.shape
    if ov.size == 0:
        return
    mx = np.max(np.abs(ov))
    big = np.abs(ov) > 1e-3 * mx
    rel = np.abs(v[big] - ov[big]) / np.abs(ov[big])
    assert rel.size == 0 or np.max(rel) <= rtol, ("relative", float(np.max(rel)), int(np.argmax(rel)))
    assert np.max(np.abs(v[~big] - ov[~big]), initial=0.0) <= rtol * mx, "absolute"


def _operator(fb2, N, p, ncomp, kind, func, prm, mat, parts=(1, 1, 1), part=0, scatter=0):
    s = fb2.FeSpace(N, order=p, num_components=ncomp, num_parts_x_dim=parts, part_id=part)
    op = fb2.FeAffineOperator()
    op.create(fb2.csr_format, [False], [True], [1], s, fb2.DiscreteIntegration(kind, func, prm, mat, scatter))
    op.compute()
    return s, op


@pytest.fixture(params=[0, 2], ids=["asm_default", "asm_scatter_only"])
def asm_variant(request, fb2):
    fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(request.param))
    yield request.param
    fb2._lib.check(fb2._lib.lib().fb2_set_assembly_variant(0))


@pytest.mark.parametrize("N,p,func,prm", [((32, 32, 32), 2, FUNC_LINEAR, None), ((40, 36, 33), 2, FUNC_SHARP_WAVE, WAVE),
                                          ((48, 48, 48), 1, FUNC_SHARP_WAVE, WAVE), ((20, 18, 16), 3, FUNC_LINEAR, None)])
def test_midsize_serial_entrywise(fb2, asm_variant, N, p, func, prm):
    s, op = _operator(fb2, N, p, 1, fb2.poisson, func, prm, None)
    o = OracleSpace(N, order=p)
    o.assemble(KIND_POISSON, func, prm, route=1)
    A = op.get_matrix()
    irp, ja = A.get_pattern()
    assert np.array_equal(irp - 1, o.irp()) and np.array_equal(ja, o.ja())
    entrywise_close(A.get_values(), o.val())
    b, ob = op.get_translation().get_entries(), o.rhs()
    assert np.max(np.abs(b - ob)) <= 1e-12 * max(np.max(np.abs(ob)), 1e-300)
    # force_compute into the same pattern is idempotent
    v1 = A.get_values()
    op.force_compute()
    entrywise_close(A.get_values(), o.val())
    assert np.max(np.abs(A.get_values() - v1)) <= 1e-13 * np.max(np.abs(v1))
    # one SpMV, every kernel variant bit-identical to the default
    x = np.random.default_rng(7).standard_normal(s.num_free_dofs)
    vx, vy = s.create_dof_values(), s.create_dof_values()
    vx.set_entries(x)
    A.apply(vx, vy)
    y0 = vy.get_entries()
    oy = o.spmv(x)
    assert np.max(np.abs(y0 - oy)) <= 1e-12 * np.max(np.abs(oy))
    # 25 Jacobi-PCG iterations: residual history against the oracle
    M = fb2.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    ils = fb2.IterativeLinearSolver(); ils.set_type_from_string("CG")
    ils.set_parameters_from_pl({fb2.ILS_RTOL: 0.0, fb2.ILS_ATOL: -1.0, fb2.ILS_MAX_NUM_ITERATIONS: 25,
                                fb2.ILS_TRACK_CONVERGENCE_HISTORY: True})
    ils.set_operators(A, M)
    xs = s.create_dof_values()
    ils.apply(op.get_translation(), xs)
    o.jacobi_setup()
    r = o.cg(prec=1, rtol=0.0, atol=-1.0, maxit=25, history=True)
    assert ils.get_num_iterations() == r["iters"] == 25
    assert np.max(np.abs(ils.history - r["history"]) / r["history"]) <= 1e-8
    assert np.max(np.abs(xs.get_entries() - r["x"])) <= 1e-9 * np.max(np.abs(r["x"]))


def test_midsize_spmv_variants_bitwise(fb2, monkeypatch):
    """the same values (the atomic scatter of the boundary rows is not order-deterministic, so they are copied) through every
    SpMV kernel variant and both row maps of the consumer lane groups (natural / transposed with staggered chunks): class-table
    columns must reproduce the ja stream bit for bit, and the row map must not change a bit of y"""
    import ctypes as C
    N, p = (32, 32, 32), 2
    s, op = _operator(fb2, N, p, 1, fb2.poisson, FUNC_SHARP_WAVE, WAVE, None)
    v = op.get_matrix().get_values()
    x = np.random.default_rng(3).standard_normal(s.num_free_dofs)
    L = fb2._lib.lib()
    ys, seen_tr = {}, set()
    try:
        for variant, tr in ((0, "0"), (0, "1"), (3, "0"), (3, "1"), (0, None), (1, None)):   # 3 = ja streamed for every tile
            if tr is None:
                monkeypatch.delenv("FB2_SPMV_TRANSPOSE", raising=False)
            else:
                monkeypatch.setenv("FB2_SPMV_TRANSPOSE", tr)
            fb2._lib.check(L.fb2_set_spmv_variant(variant))
            A2 = fb2.SparseMatrix(s)                   # the variant binds when the matrix is first applied
            A2.set_values(v)
            vx2, vy2 = s.create_dof_values(), s.create_dof_values()
            vx2.set_entries(x)
            A2.apply(vx2, vy2)
            ys[(variant, tr)] = vy2.get_entries()
            A2.apply(vx2, vy2)
            assert np.array_equal(ys[(variant, tr)], vy2.get_entries()), "SpMV is not run-to-run deterministic"
            if variant != 1:
                ms = C.c_double(0)
                fb2._lib.check(L.fb2_last_phase_ms(b"spmv_transposed_rows", C.byref(ms)))
                seen_tr.add(int(ms.value))
                if tr is not None:
                    assert int(ms.value) == int(tr)
    finally:
        fb2._lib.check(L.fb2_set_spmv_variant(0))
    assert seen_tr == {0, 1}
    ref = ys[(3, "0")]
    for key in ((0, "0"), (0, "1"), (3, "1"), (0, None)):
        assert np.array_equal(ys[key], ref), ("SpMV differs bitwise from the natural-map ja-streaming SpMV", key)
    assert np.max(np.abs(ref - ys[(1, None)])) <= 1e-13 * np.max(np.abs(ys[(1, None)]))


@pytest.mark.parametrize("N,p,parts", [((32, 32, 32), 2, (2, 2, 2)), ((24, 24, 24), 1, (2, 2, 1))])
def test_midsize_par_local_matrices_entrywise(fb2, asm_variant, N, p, parts):
    for part in range(1, int(np.prod(parts)) + 1):
        s, op = _operator(fb2, N, p, 1, fb2.poisson, FUNC_LINEAR, None, None, parts, part)
        o = OracleSpace(N, order=p, parts=list(parts), part_id=part)
        o.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
        irp, ja = op.get_matrix().get_pattern()
        assert np.array_equal(irp - 1, o.irp()) and np.array_equal(ja, o.ja())
        entrywise_close(op.get_matrix().get_values(), o.val())
        b, ob = op.get_translation().get_entries(), o.rhs()
        assert np.max(np.abs(b - ob)) <= 1e-12 * max(np.max(np.abs(ob)), 1e-300)
        x = np.random.default_rng(part).standard_normal(s.num_free_dofs)
        vx, vy = s.create_dof_values(), s.create_dof_values()
        vx.set_entries(x)
        op.get_matrix().apply(vx, vy)                      # no communicator in this process: the local matvec
        oy = o.spmv(x)
        assert np.max(np.abs(vy.get_entries() - oy)) <= 1e-12 * np.max(np.abs(oy))
        op.free(); s.free(); o.close()


@pytest.mark.parametrize("N,p,kind,okind,mat", [((8, 8, 8), 2, 1, KIND_VECTOR_LAPLACIAN, None),
                                                ((9, 8, 8), 2, 2, KIND_ELASTICITY, [5.0 / 18.0, 5.0 / 12.0]),
                                                ((12, 12, 12), 1, 2, KIND_ELASTICITY, [5.0 / 18.0, 5.0 / 12.0]),
                                                ((12, 11, 10), 1, 1, KIND_VECTOR_LAPLACIAN, None)])
def test_midsize_vector_operators_entrywise(fb2, asm_variant, N, p, kind, okind, mat):
    s, op = _operator(fb2, N, p, 3, kind, FUNC_LINEAR, None, mat)
    o = OracleSpace(N, order=p, ncomp=3)
    o.assemble(okind, FUNC_LINEAR, mat=mat, route=1)
    irp, ja = op.get_matrix().get_pattern()
    assert np.array_equal(irp - 1, o.irp()) and np.array_equal(ja, o.ja())
    entrywise_close(op.get_matrix().get_values(), o.val())
    b, ob = op.get_translation().get_entries(), o.rhs()
    assert np.max(np.abs(b - ob)) <= 1e-12 * max(np.max(np.abs(ob)), 1e-300)
