"""torchrun worker: the reference's 8-part (2x2x2) layout on FEWER GPUs than parts (BASELINE config 3 on 2 or 4 GPUs): every
rank holds 8/world co-resident sub-domains; comm(), weighted dots, Jacobi set-up and PCG run collectively over them and are
checked against the in-process 8-part oracle."""
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
    parts = (2, 2, 2)
    nparts = 8
    k = nparts // world
    fb.comm_set_parts_per_rank(k)
    mine = list(range(rank * k + 1, rank * k + k + 1))            # 1-based part ids of this rank
    ok, fails = True, []

    def chk(name, cond, info=""):
        nonlocal ok
        if not cond:
            ok = False
            fails.append("rank %d %s %s" % (rank, name, info))

    for N, p, func, prm, prec, stop in [((8, 8, 8), 2, FUNC_SHARP_WAVE, WAVE, 1, "RES_RES"), ((8, 8, 8), 1, FUNC_LINEAR, None, 0, "RES_RES"),
                                        ((6, 6, 6), 2, FUNC_SHARP_WAVE, WAVE, 1, "DELTA_RHS_AND_RES_RES")]:
        P = ParOracle(N, order=p, parts=parts)
        P.assemble(KIND_POISSON, func, prm)
        P.jacobi_setup()
        stop_id = {"RES_RES": 0, "DELTA_RHS_AND_RES_RES": 6}[stop]
        ro = P.cg(prec=prec, rtol=1e-12, maxit=5000, stop=stop_id)
        spaces = [fb.FeSpace(N, order=p, num_parts_x_dim=parts, part_id=q) for q in mine]
        ops = []
        for s in spaces:
            op = fb.FeAffineOperator()
            op.create(fb.csr_format, [False], [True], [1], s, fb.DiscreteIntegration(fb.poisson, func, prm))
            op.compute()                                           # local (sub-assembled) matrix and RHS, no communication
            ops.append(op)
        As = [op.get_matrix() for op in ops]
        bs = [op.get_translation() for op in ops]
        fb.vec_comm_multi(bs)                                      # par_sparse_assembler: RHS comm()
        for s, q, bv in zip(spaces, mine, bs):
            ob = P.b[q - 1]
            chk("rhs", np.max(np.abs(bv.get_entries() - ob)) <= 1e-12 * max(np.max(np.abs(ob)), 1e-300), q)
        ones = [s.create_dof_values() for s in spaces]
        for v in ones:
            v.init(1.0)
        fb.vec_comm_multi(ones)
        oo = P.comm([np.ones(sp.num_free) for sp in P.spaces])
        for q, v in zip(mine, ones):
            chk("comm_ones", np.array_equal(v.get_entries(), oo[q - 1]), q)
        xs_o = P.comm([np.random.default_rng(7 + i).standard_normal(sp.num_free) * im["weight"] for i, (sp, im) in
                       enumerate(zip(P.spaces, P.imp))])
        vxs, vys = [s.create_dof_values() for s in spaces], [s.create_dof_values() for s in spaces]
        for q, v in zip(mine, vxs):
            v.set_entries(xs_o[q - 1])
        fb.spmv_multi(As, vxs, vys)
        oy = P.spmv(xs_o)
        for q, v in zip(mine, vys):
            chk("spmv", np.max(np.abs(v.get_entries() - oy[q - 1])) <= 1e-12 * np.max(np.abs(oy[q - 1])), q)
        od = P.dot(xs_o, oy)
        chk("dot", abs(fb.vec_dot_multi(vxs, vys) - od) <= 1e-11 * abs(od), (fb.vec_dot_multi(vxs, vys), od))
        dinvs = None
        if prec == 1:
            dinvs = [s.create_dof_values() for s in spaces]
            fb.jacobi_setup_multi(As, dinvs)
            for q, d in zip(mine, dinvs):
                chk("dinv", np.max(np.abs(d.get_entries() - P.dinv[q - 1])) <= 1e-12 * np.max(np.abs(P.dinv[q - 1])), q)
        ils = fb.IterativeLinearSolver()
        ils.set_parameters_from_pl({fb.ILS_TYPE: "CG", fb.ILS_RTOL: 1e-12, fb.ILS_MAX_NUM_ITERATIONS: 5000,
                                    fb.ILS_STOPPING_CRITERIUM: stop, fb.ILS_TRACK_CONVERGENCE_HISTORY: True})
        sols = [s.create_dof_values() for s in spaces]
        ils.apply_multi(As, dinvs, bs, sols)
        chk("iters", ils.converged() and abs(ils.get_num_iterations() - ro["iters"]) <= 1, (ils.get_num_iterations(), ro["iters"]))
        kk = min(ils.get_num_iterations(), ro["iters"]) // 2
        chk("history", np.allclose(ils.history[:kk], ro["history"][:kk], rtol=1e-6))
        for q, v in zip(mine, sols):
            chk("solution", np.max(np.abs(v.get_entries() - ro["x"][q - 1])) <= 1e-8, q)
        if rank == 0:
            print("case", N, p, func, prec, stop, "parts/rank", k, "iters", ils.get_num_iterations(), ro["iters"], "ok", ok, flush=True)
    for f in fails:
        print("FAIL", f, flush=True)
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0 and int(t.item()) == 1:
        print("PAR_MULTI_OK")
    fb.comm_finalize()
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
