"""torchrun worker of tests/test_gpu_par.py::test_par_pcg_nccl: one rank = one GPU = one FEMPAR sub-domain.
Checks the NCCL path (RHS comm, halo exchange in SpMV, weighted dots + all-reduce, Jacobi set-up, PCG) of this rank's part
against the in-process multi-part oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import fempar_b200 as fb  # noqa: E402
from oracle.oracle import KIND_POISSON, FUNC_LINEAR, FUNC_SHARP_WAVE  # noqa: E402
from oracle.par_oracle import ParOracle  # noqa: E402

WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    fb.init(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fb.comm_init_from_torch(dist)
    parts = {2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[world]
    ok = True
    fails = []

    def chk(name, cond, info=""):
        nonlocal ok
        if not cond:
            ok = False
            fails.append("rank %d %s %s" % (rank, name, info))

    for N, p, func, prm, prec in [((8, 8, 8), 1, FUNC_LINEAR, None, 1), ((8, 8, 8), 2, FUNC_SHARP_WAVE, WAVE, 1),
                                  ((8, 8, 8), 2, FUNC_LINEAR, None, 0), ((4, 4, 4), 3, FUNC_LINEAR, None, 1)]:
        P = ParOracle(N, order=p, parts=parts)
        P.assemble(KIND_POISSON, func, prm)
        P.jacobi_setup()
        ro = P.cg(prec=prec, rtol=1e-12, maxit=5000)
        s = fb.FeSpace(N, order=p, num_parts_x_dim=parts, part_id=rank + 1)
        op = fb.FeAffineOperator()
        op.create(fb.csr_format, [False], [True], [1], s, fb.DiscreteIntegration(fb.poisson, func, prm))
        op.compute()
        b = op.get_translation().get_entries()
        ob = P.b[rank]
        chk("rhs", np.max(np.abs(b - ob)) <= 1e-12 * max(np.max(np.abs(ob)), 1e-300), np.max(np.abs(b - ob)))   # RHS after comm()
        ones = s.create_dof_values(); ones.init(1.0); ones.comm()
        oo = P.comm([np.ones(q.num_free) for q in P.spaces])[rank]
        chk("comm_ones", np.array_equal(ones.get_entries(), oo))
        x = np.random.default_rng(7).standard_normal(s.num_free_dofs)
        vx, vy = s.create_dof_values(), s.create_dof_values()
        # a consistent ("fully assembled") input: make interface copies agree through comm of owner values
        vx.set_entries(x * s.get_dof_import()["weight"]); vx.comm()
        xs = P.comm([np.random.default_rng(7).standard_normal(q.num_free) * i["weight"] for q, i in zip(P.spaces, P.imp)])
        op.get_matrix().apply(vx, vy)
        oy = P.spmv(xs)[rank]
        chk("spmv", np.max(np.abs(vy.get_entries() - oy)) <= 1e-12 * np.max(np.abs(oy)), np.max(np.abs(vy.get_entries() - oy)) / np.max(np.abs(oy)))
        od = P.dot(xs, P.spmv(xs))
        chk("dot", abs(vx.dot(vy) - od) <= 1e-11 * abs(od), (vx.dot(vy), od))
        M = None
        if prec == 1:
            M = fb.JacobiPreconditioner(); M.create(op); M.numerical_setup()
            chk("dinv", np.max(np.abs(M.inverse_diagonal.get_entries() - P.dinv[rank])) <= 1e-12 * np.max(np.abs(P.dinv[rank])))
        ils = fb.IterativeLinearSolver(); ils.set_type_from_string("CG")
        ils.set_parameters_from_pl({fb.ILS_RTOL: 1e-12, fb.ILS_MAX_NUM_ITERATIONS: 5000})
        ils.set_operators(op.get_tangent(), M)
        sol = s.create_dof_values()
        ils.apply(op.get_translation(), sol)
        chk("iters", ils.converged() and abs(ils.get_num_iterations() - ro["iters"]) <= 1, (ils.get_num_iterations(), ro["iters"]))
        chk("solution", np.max(np.abs(sol.get_entries() - ro["x"][rank])) <= 1e-8, np.max(np.abs(sol.get_entries() - ro["x"][rank])))
        h1, l2 = s.error_norms(sol, func, prm)
        oh1, ol2 = P.errors(ro["x"], func, prm)
        chk("errors", abs(h1 - oh1) <= 1e-8 * max(oh1, 1.0) and abs(l2 - ol2) <= 1e-8 * max(ol2, 1.0), (h1, oh1, l2, ol2))
        if rank == 0:
            print("case", N, p, func, prec, "iters", ils.get_num_iterations(), ro["iters"], "ok", ok, flush=True)
    # --- host-built objects (SURVEY 8b comm / matrix-format seams): the oracle's dof_import_t and local CSR matrix are
    #     handed over through fb2_dof_import_set / fb2_csr_create_from_host; the par PCG must give the same answer ---
    N, p = (8, 8, 8), 2
    P = ParOracle(N, order=p, parts=parts)
    P.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE)
    P.jacobi_setup()
    ro = P.cg(prec=1, rtol=1e-12, maxit=5000)
    q, im = P.spaces[rank], P.imp[rank]
    bs = fb.BareFeSpace(q.num_free)
    bs.set_dof_import(im["list_rcv"], im["rcv_ptrs"] + 1, im["unpack_idx"], im["list_snd"], im["snd_ptrs"] + 1, im["pack_idx"])
    Ah = fb.SparseMatrix(space=bs, host_csr=(q.irp() + 1, q.ja(), q.val()))
    Mh = fb.JacobiPreconditioner(); Mh.create(Ah); Mh.numerical_setup()
    chk("host_dinv", np.max(np.abs(Mh.inverse_diagonal.get_entries() - P.dinv[rank])) <= 1e-12 * np.max(np.abs(P.dinv[rank])))
    bh = bs.create_dof_values(); bh.set_entries(P.b[rank])
    xh = bs.create_dof_values()
    ils = fb.IterativeLinearSolver(); ils.set_type_from_string("CG")
    ils.set_parameters_from_pl({fb.ILS_RTOL: 1e-12, fb.ILS_MAX_NUM_ITERATIONS: 5000})
    ils.set_operators(Ah, Mh)
    ils.apply(bh, xh)
    chk("host_iters", ils.converged() and abs(ils.get_num_iterations() - ro["iters"]) <= 1, (ils.get_num_iterations(), ro["iters"]))
    chk("host_solution", np.max(np.abs(xh.get_entries() - ro["x"][rank])) <= 1e-8)
    if rank == 0:
        print("host-built import + CSR: iters", ils.get_num_iterations(), ro["iters"], "ok", ok, flush=True)
    # --- host-numbered par space: the oracle's per-part lst_dofs_gids (local + ghost cells, interior-then-
    #     interface numbering), vertex coordinates by global cell id and dof_import_t are uploaded; affine coordinates must
    #     reproduce the par oracle, warped coordinates must pass the patch test across the parts ---
    from oracle import hex_oracle as ho
    cg_ids = np.asarray(q.cells_gid()) - 1
    warp = lambda X: X + 0.05 * np.sin(np.pi * X) * np.sin(np.pi * X[..., [1, 2, 0]])
    for tag, wfun in (("affine", None), ("warped", warp)):
        Xc = ho.structured_cell_coords(N[0], wfun)[cg_ids]
        gl = q.lst_dofs_gids().copy()
        hs = fb.HostNumberedFeSpace(p, 1, gl, Xc, q.num_free, q.num_dirichlet, num_local_cells=q.num_local_cells)
        hs.set_dof_import(im["list_rcv"], im["rcv_ptrs"] + 1, im["unpack_idx"], im["list_snd"], im["snd_ptrs"] + 1, im["pack_idx"])
        func, prm = (FUNC_SHARP_WAVE, WAVE) if wfun is None else (FUNC_LINEAR, None)
        hop = fb.FeAffineOperator()
        hop.create(fb.csr_format, [False], [True], [1], hs, fb.DiscreteIntegration(fb.poisson, func, prm))
        hop.compute()
        hM = fb.JacobiPreconditioner(); hM.create(hop); hM.numerical_setup()
        ils = fb.IterativeLinearSolver(); ils.set_type_from_string("CG")
        ils.set_parameters_from_pl({fb.ILS_RTOL: 1e-12, fb.ILS_MAX_NUM_ITERATIONS: 5000})
        ils.set_operators(hop.get_tangent(), hM)
        hx = hs.create_dof_values()
        ils.apply(hop.get_translation(), hx)
        if wfun is None:
            hb = hop.get_translation().get_entries()
            chk("hostnum_rhs", np.max(np.abs(hb - P.b[rank])) <= 1e-12 * np.max(np.abs(P.b[rank])))
            chk("hostnum_dinv", np.max(np.abs(hM.inverse_diagonal.get_entries() - P.dinv[rank])) <= 1e-12 * np.max(np.abs(P.dinv[rank])))
            chk("hostnum_iters", ils.converged() and abs(ils.get_num_iterations() - ro["iters"]) <= 1, (ils.get_num_iterations(), ro["iters"]))
            chk("hostnum_solution", np.max(np.abs(hx.get_entries() - ro["x"][rank])) <= 1e-8)
        else:
            xn = ho.node_coordinates(p, Xc).sum(-1)
            exact = np.zeros(q.num_free)
            m = (gl > 0) & (gl <= q.num_free)
            exact[gl[m] - 1] = xn[m]
            chk("hostnum_patch", ils.converged() and np.max(np.abs(hx.get_entries() - exact)) <= 1e-9, np.max(np.abs(hx.get_entries() - exact)))
            h1, l2 = hs.error_norms(hx, FUNC_LINEAR)
            chk("hostnum_errors", h1 < 1e-8 and l2 < 1e-9, (h1, l2))
        if rank == 0:
            print("host-numbered par space", tag, "iters", ils.get_num_iterations(), "ok", ok, flush=True)
        hop.free()
    # --- other Krylov solvers in the par layout: iteration counts and residual histories are invariant
    #     under the DOF renumbering of the partition, so the SERIAL restatement on the same mesh is the checker ---
    import scipy.sparse as sp
    from oracle.oracle import OracleSpace
    from oracle import krylov_oracle as K
    so = OracleSpace(N, order=p)
    so.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    sd = so.jacobi_setup().copy()
    SA = sp.csr_matrix((so.val().copy(), (so.ja() - 1).copy(), so.irp().copy()), shape=(so.num_free, so.num_free))
    sb = so.rhs().copy()
    s = fb.FeSpace(N, order=p, num_parts_x_dim=parts, part_id=rank + 1)
    op = fb.FeAffineOperator()
    op.create(fb.csr_format, [False], [True], [1], s, fb.DiscreteIntegration(fb.poisson, FUNC_SHARP_WAVE, WAVE))
    op.compute()
    M = fb.JacobiPreconditioner(); M.create(op); M.numerical_setup()
    for name, fn, kw in [("RGMRES", K.rgmres, dict(dkrymax=12)), ("FGMRES", K.fgmres, dict(dkrymax=40, strat=K.MGSRO)),
                         ("LGMRES", K.lgmres, dict(dkrymax=12)), ("ICG", K.icg, {}),
                         ("RICHARDSON", K.richardson, dict(relaxation=0.8))]:
        rt, mx = (1e-2, 300) if name == "RICHARDSON" else (1e-9, 1000)
        ref = fn(lambda v: SA @ v, lambda r: sd * r, sb, rtol=rt, maxit=mx, **kw)
        ils = fb.IterativeLinearSolver()
        pl = {fb.ILS_TYPE: name, fb.ILS_RTOL: rt, fb.ILS_MAX_NUM_ITERATIONS: mx, fb.ILS_TRACK_CONVERGENCE_HISTORY: True}
        if "dkrymax" in kw:
            pl[fb.ILS_MAX_DIM_KRYLOV_BASIS] = kw["dkrymax"]
            pl[fb.ILS_ORTHONORM_STRATEGY] = "MGSRO" if kw.get("strat") == K.MGSRO else "ICGSRO"
        if "relaxation" in kw:
            pl[fb.ILS_RELAXATION] = kw["relaxation"]
        ils.set_parameters_from_pl(pl)
        ils.set_operators(op.get_tangent(), M)
        sol = s.create_dof_values()
        ils.apply(op.get_translation(), sol)
        kk = min(ils.get_num_iterations(), ref["iters"]) * 2 // 3
        chk(name + "_iters", ils.converged() == ref["converged"] and abs(ils.get_num_iterations() - ref["iters"]) <= 1,
            (ils.get_num_iterations(), ref["iters"]))
        chk(name + "_history", np.allclose(ils.history[:kk], ref["history"][:kk], rtol=1e-5))
        h1, l2 = s.error_norms(sol, FUNC_SHARP_WAVE, WAVE)
        oh1, ol2 = so.errors(ref["x"], FUNC_SHARP_WAVE, WAVE)
        chk(name + "_errors", abs(h1 - oh1) <= 1e-5 * oh1 and abs(l2 - ol2) <= 1e-5 * ol2, (h1, oh1, l2, ol2))
        if rank == 0:
            print("par", name, "iters", ils.get_num_iterations(), ref["iters"], "ok", ok, flush=True)
    for f in fails:
        print("FAIL", f, flush=True)
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0 and int(t.item()) == 1:
        print("PAR_NCCL_OK")
    fb.comm_finalize()
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
