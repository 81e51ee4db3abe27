"""World-size-2 gloo worker (CPU): the host-side protocol of the N>1 path -- one rank per sub-domain, neighbour exchange by
point-to-point messages in dof_import list order (mpi_context.f90:522-639), dots by all-reduce (mpi_context.f90:384) --
run on the oracle's per-part data and checked against the single-process multi-part oracle."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR  # noqa: E402
from oracle.par_oracle import ParOracle  # noqa: E402


def exchange(x, imp, reverse, add):
    snd = (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"]) if reverse else (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"])
    rcv = (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"]) if reverse else (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"])
    reqs, bufs = [], []
    for i, q in enumerate(rcv[0]):
        b = torch.empty(int(rcv[1][i + 1] - rcv[1][i]), dtype=torch.float64)
        bufs.append(b)
        reqs.append(dist.irecv(b, src=int(q) - 1))
    for i, q in enumerate(snd[0]):
        reqs.append(dist.isend(torch.from_numpy(x[snd[2][snd[1][i]:snd[1][i + 1]] - 1].copy()), dst=int(q) - 1))
    for r in reqs:
        r.wait()
    for i, b in enumerate(bufs):
        ii = rcv[2][rcv[1][i]:rcv[1][i + 1]] - 1
        if add:
            np.add.at(x, ii, b.numpy())
        else:
            x[ii] = b.numpy()


def comm(x, imp):
    exchange(x, imp, False, True)
    exchange(x, imp, True, False)
    return x


def pdot(x, y, imp):
    t = torch.tensor([float(np.dot(imp["weight"] * x, y))], dtype=torch.float64)
    dist.all_reduce(t)
    return float(t.item())


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    parts = {2: (2, 1, 1), 4: (2, 2, 1)}[world]
    N, p = (6, 6, 6), 2
    s = OracleSpace(N, order=p, parts=list(parts), part_id=rank + 1)
    imp = s.dof_import()
    s.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
    b = comm(s.rhs().copy(), imp)
    dinv = 1.0 / comm(1.0 / s.jacobi_setup(), imp)
    # PCG, cg.f90:101-238
    x = np.zeros(s.num_free)
    r = b - comm(s.spmv(x), imp)
    z = dinv * r
    rz = pdot(r, z, imp)
    pv = z.copy()
    tol = 1e-12 * np.sqrt(pdot(r, r, imp))
    it = 0
    while True:
        it += 1
        Ap = comm(s.spmv(pv), imp)
        App = pdot(Ap, pv, imp)
        alpha = rz / App if App != 0.0 else 0.0
        x += alpha * pv
        r -= alpha * Ap
        e = np.sqrt(pdot(r, r, imp))
        if e <= tol or it >= 5000:
            break
        z = dinv * r
        rz_new = pdot(r, z, imp)
        pv = z + (rz_new / rz) * pv
        rz = rz_new
    P = ParOracle(N, order=p, parts=parts)
    P.assemble(KIND_POISSON, FUNC_LINEAR)
    P.jacobi_setup()
    ro = P.cg(prec=1, rtol=1e-12, maxit=5000)
    ok = abs(it - ro["iters"]) <= 1 and np.max(np.abs(x - ro["x"][rank])) < 1e-9 and np.max(np.abs(b - P.b[rank])) < 1e-13
    t = torch.tensor([1 if ok else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("iters", it, ro["iters"])
        if int(t.item()) == 1:
            print("PAR_GLOO_OK")
    dist.destroy_process_group()
    sys.exit(0 if int(t.item()) == 1 else 1)


if __name__ == "__main__":
    main()
