"""CPU tests of the oracle's par_* layer (dof_import, comm, weighted dot, par CG) -- the checker of the multi-GPU path.

Acceptance = the reference's own par_test_poisson criterion (Tests/Par/par_test_poisson/par_test_poisson_driver.f90:557-589:
u = x+y+z, CG rtol 1e-12, all error norms < 1e-4) on the partitions of Tests/Scripts/par_test_poisson_fast:23-67,
with Jacobi and identity preconditioners; plus structural invariants of dof_import_t.
"""
import numpy as np
import pytest

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR
from oracle.par_oracle import ParOracle


@pytest.mark.parametrize("N,p,parts", [((8, 8, 8), 1, (2, 2, 1)), ((12, 12, 12), 1, (3, 3, 3)), ((8, 8, 8), 2, (2, 2, 2)),
                                       ((6, 6, 6), 3, (2, 1, 3)), ((4, 4, 4), 2, (4, 2, 1))])
@pytest.mark.parametrize("prec", [0, 1])
def test_par_test_poisson_acceptance(N, p, parts, prec):
    P = ParOracle(N, order=p, parts=parts)
    P.assemble(KIND_POISSON, FUNC_LINEAR)
    P.jacobi_setup()
    r = P.cg(prec=prec, rtol=1e-12, maxit=5000)
    assert r["converged"]
    h1, l2 = P.errors(r["x"])
    assert h1 < 1e-4 and l2 < 1e-4            # the reference's bar
    assert h1 < 1e-9 and l2 < 1e-9            # what exact representability of x+y+z really gives
    s = OracleSpace(N, order=p)
    s.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
    s.jacobi_setup()
    rs = s.cg(prec=prec, rtol=1e-12, maxit=5000)
    assert abs(r["iters"] - rs["iters"]) <= 1   # par CG is algebraically the serial CG


@pytest.mark.parametrize("N,p,parts", [((4, 4, 4), 1, (2, 1, 1)), ((6, 6, 6), 2, (2, 3, 1)), ((6, 6, 6), 1, (3, 3, 3)), ((4, 4, 4), 3, (2, 2, 2))])
def test_dof_import_invariants(N, p, parts):
    P = ParOracle(N, order=p, parts=parts)
    n_glob = int(np.prod([p * n - 1 for n in N]))
    assert sum(int(i["weight"].sum()) for i in P.imp) == n_glob      # every global DOF owned exactly once
    for q, (s, imp) in enumerate(zip(P.spaces, P.imp)):
        nI, nG = s.num_interior, s.num_interface
        assert nI + nG == s.num_free
        for idx in (imp["pack_idx"], imp["unpack_idx"]):               # interface DOFs are numbered last (dof_import.f90:391-421)
            assert idx.size == 0 or (idx.min() > nI and idx.max() <= s.num_free)
        assert np.all(imp["weight"][:nI] == 1.0)
        assert np.all(np.diff(imp["list_rcv"]) > 0) and np.all(np.diff(imp["list_snd"]) > 0)
        # what I send to r is exactly what r expects from me
        for i, r in enumerate(imp["list_snd"]):
            o = P.imp[r - 1]
            k = list(o["list_rcv"]).index(q + 1)
            assert imp["snd_ptrs"][i + 1] - imp["snd_ptrs"][i] == o["rcv_ptrs"][k + 1] - o["rcv_ptrs"][k]
    # comm() of a vector of ones counts the parts sharing each DOF: 1 interior, 2 face, 4 edge, 8 corner
    ones = P.comm([np.ones(s.num_free) for s in P.spaces])
    for v in ones:
        assert set(np.unique(v)).issubset({1.0, 2.0, 4.0, 8.0})


def test_two_part_split_is_half_half():
    # sbm_par_fe_space.i90:2807-2832: couplings shared by exactly two parts are split half/half, the max part owns the first half
    P = ParOracle((4, 4, 4), order=1, parts=(2, 1, 1))
    a, b = P.imp
    assert a["pack_idx"].size == 4 and a["unpack_idx"].size == 5
    assert b["unpack_idx"].size == 4 and b["pack_idx"].size == 5


@pytest.mark.parametrize("N,p,parts", [((8, 8, 8), 2, (2, 2, 2)), ((8, 8, 8), 1, (4, 2, 1)), ((6, 6, 6), 2, (1, 1, 1))])
def test_threads_as_ranks_run_matches_numpy_par_oracle(N, p, parts):
    """fo_par_run (the timed MPI-layout CPU baseline of bench.py: one host thread per rank, mailbox exchange, weighted dots)
    leaves the same |b|, |r| after 25 PCG iterations and assembled diagonal as the numpy restatement of the par layer."""
    from oracle.par_oracle import par_run_threads
    nparts = int(np.prod(parts))
    if nparts == 1:
        s = OracleSpace(N, order=p)
        r = par_run_threads([s], cg_iters=25)
        s.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
        b = s.rhs().copy()
        dinv = s.jacobi_setup()
        c = s.cg(prec=1, rtol=0.0, atol=-1.0, maxit=25)
        res = b - s.spmv(c["x"])
        ref = (np.linalg.norm(b), np.linalg.norm(res), float(np.sum(1.0 / dinv)))
    else:
        P = ParOracle(N, order=p, parts=parts)
        r = par_run_threads(P.spaces, cg_iters=25)
        b = P.assemble(KIND_POISSON, FUNC_LINEAR)
        P.jacobi_setup()
        c = P.cg(prec=1, rtol=0.0, atol=-1.0, maxit=25)
        Ax = P.spmv([v.copy() for v in c["x"]])
        res = [bv - av for bv, av in zip(b, Ax)]
        ref = (np.sqrt(P.dot(b, b)), np.sqrt(P.dot(res, res)),
               float(sum(np.dot(i["weight"], 1.0 / d) for i, d in zip(P.imp, P.dinv))))
    assert abs(r["b_nrm2"] - ref[0]) <= 1e-12 * ref[0]
    assert abs(r["r_nrm2"] - ref[1]) <= 1e-8 * max(ref[1], 1e-12) + 1e-14
    assert abs(r["diag_sum"] - ref[2]) <= 1e-12 * ref[2]
    assert r["assembly_s"] > 0 and r["apply_s"] > 0 and r["pcg_s"] > 0
