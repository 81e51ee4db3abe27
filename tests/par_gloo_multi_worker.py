"""World-size-2 gloo worker (CPU): the host-side protocol of SEVERAL sub-domains per process (the 8-part layout on 2 ranks):
co-resident parts exchange by local copies, remote ones by point-to-point messages that both sides order by
(source part, destination part) -- gloo, like NCCL, matches the messages of a rank pair in issue order.  Mirrors
exchange_stage_multi of fempar_b200/csrc/fb2_comm.cu on the oracle's per-part data."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_SHARP_WAVE  # noqa: E402
from oracle.par_oracle import ParOracle  # noqa: E402

WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def exchange_multi(xs, imps, mine, k, rank, reverse, add):
    rank_of = lambda part: (part - 1) // k
    sends, recvs, rbuf = [], [], {}
    for i, (x, imp) in enumerate(zip(xs, imps)):
        sl, sp, si = (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"]) if reverse else (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"])
        rl, rp, _ = (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"]) if reverse else (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"])
        for a, nb in enumerate(sl):
            data = x[si[sp[a]:sp[a + 1]] - 1].copy()
            if rank_of(int(nb)) == rank:
                rbuf[(mine[i], int(nb))] = data                      # local "copy": (src, dst) -> payload
            else:
                sends.append((mine[i], int(nb), data))
        for a, nb in enumerate(rl):
            if rank_of(int(nb)) != rank:
                recvs.append((int(nb), mine[i], int(rp[a + 1] - rp[a])))
    sends.sort(key=lambda m: (m[0], m[1]))
    recvs.sort(key=lambda m: (m[0], m[1]))
    reqs = []
    for src, dst, n in recvs:
        b = torch.empty(n, dtype=torch.float64)
        rbuf[(src, dst)] = b
        reqs.append(dist.irecv(b, src=rank_of(src)))
    for src, dst, data in sends:
        reqs.append(dist.isend(torch.from_numpy(data), dst=rank_of(dst)))
    for r in reqs:
        r.wait()
    for i, (x, imp) in enumerate(zip(xs, imps)):
        rl, rp, ri = (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"]) if reverse else (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"])
        for a, nb in enumerate(rl):                                  # neighbour-list order = the reference's order of additions
            b = rbuf[(int(nb), mine[i])]
            b = b.numpy() if isinstance(b, torch.Tensor) else b
            ii = ri[rp[a]:rp[a + 1]] - 1
            if add:
                np.add.at(x, ii, b)
            else:
                x[ii] = b


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    parts, nparts = (2, 2, 2), 8
    k = nparts // world
    mine = list(range(rank * k + 1, rank * k + k + 1))
    N, p = (4, 4, 4), 2
    sp = [OracleSpace(N, order=p, parts=list(parts), part_id=q) for q in mine]
    imps = [s.dof_import() for s in sp]
    for s in sp:
        s.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)

    def comm(xs):
        exchange_multi(xs, imps, mine, k, rank, False, True)
        exchange_multi(xs, imps, mine, k, rank, True, False)
        return xs

    def pdot(xs, ys):
        t = torch.tensor([sum(float(np.dot(im["weight"] * x, y)) for x, y, im in zip(xs, ys, imps))], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t.item())

    b = comm([s.rhs().copy() for s in sp])
    dinv = [1.0 / d for d in comm([1.0 / s.jacobi_setup() for s in sp])]
    x = [np.zeros(s.num_free) for s in sp]
    r = [bv.copy() for bv in b]
    z = [d * rv for d, rv in zip(dinv, r)]
    rz = pdot(r, z)
    pv = [v.copy() for v in z]
    tol = 1e-12 * np.sqrt(pdot(r, r))
    it = 0
    while True:
        it += 1
        Ap = comm([s.spmv(v) for s, v in zip(sp, pv)])
        App = pdot(Ap, pv)
        alpha = rz / App if App != 0.0 else 0.0
        for xv, rv, v, av in zip(x, r, pv, Ap):
            xv += alpha * v
            rv -= alpha * av
        if np.sqrt(pdot(r, r)) <= tol or it >= 5000:
            break
        z = [d * rv for d, rv in zip(dinv, r)]
        rz_new = pdot(r, z)
        pv = [zv + (rz_new / rz) * v for zv, v in zip(z, pv)]
        rz = rz_new
    P = ParOracle(N, order=p, parts=parts)
    P.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE)
    P.jacobi_setup()
    ro = P.cg(prec=1, rtol=1e-12, maxit=5000)
    ok = abs(it - ro["iters"]) <= 1
    for q, xv, bv in zip(mine, x, b):
        ok = ok and np.max(np.abs(xv - ro["x"][q - 1])) < 1e-9 and np.max(np.abs(bv - P.b[q - 1])) < 1e-13
    t = torch.tensor([1 if ok else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("iters", it, ro["iters"])
        if int(t.item()) == 1:
            print("PAR_GLOO_MULTI_OK")
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
