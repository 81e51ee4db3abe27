"""Multi-part driver of the CPU oracle: FEMPAR's par_* layer with every part living in this process.

TEST INFRASTRUCTURE ONLY (same rule as oracle.py).  Restates, on top of the per-part data built by
fempar_oracle.cpp (dof_import lists, weights, interior-then-interface numbering):
  par_scalar_array%comm   LinearAlgebra/Vector/par_scalar_array.f90:498-533  (2-stage neighbour exchange)
  pack_rp / unpack_rp     Tools/mpi_context.f90:522-639
  par dot / nrm2          par_scalar_array.f90:303-350,451-470 (weighted local dot + l1_sum)
  par_sparse_matrix apply LinearAlgebra/SparseMatrix/par_sparse_matrix.f90:739-762 (local matvec + comm)
  jacobi numerical set-up LinearAlgebra/Preconditioner/jacobi_preconditioner.f90:356-412 (diag, comm, invert)
  cg_solve_body           LinearAlgebra/IterativeLinearSolver/cg.f90:101-238
"""
import numpy as np

from .oracle import OracleSpace, KIND_POISSON, FUNC_LINEAR


class ParOracle:
    def __init__(self, N, order=1, parts=(2, 1, 1), ncomp=1, dir_mask=0xFFFFFFFF):
        self.parts = tuple(int(p) for p in parts)
        self.nparts = int(np.prod(self.parts))
        self.spaces = [OracleSpace(N, order=order, ncomp=ncomp, parts=list(self.parts), part_id=p + 1, dir_mask=dir_mask)
                       for p in range(self.nparts)]
        self.imp = [s.dof_import() for s in self.spaces]

    # --- execution_context%neighbours_exchange with pack/unpack index lists -----------------------
    def _exchange(self, vecs, reverse, add):
        msgs = {}
        for p, (imp, x) in enumerate(zip(self.imp, vecs)):
            lst, ptr, idx = (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"]) if reverse else \
                            (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"])
            for i, q in enumerate(lst):
                msgs[(p + 1, int(q))] = x[idx[ptr[i]:ptr[i + 1]] - 1].copy()
        for p, (imp, x) in enumerate(zip(self.imp, vecs)):
            lst, ptr, idx = (imp["list_snd"], imp["snd_ptrs"], imp["pack_idx"]) if reverse else \
                            (imp["list_rcv"], imp["rcv_ptrs"], imp["unpack_idx"])
            for i, q in enumerate(lst):            # unpack neighbour by neighbour in list order
                buf = msgs[(int(q), p + 1)]
                ii = idx[ptr[i]:ptr[i + 1]] - 1
                assert buf.size == ii.size, "send/receive lists of parts %d and %d do not match" % (p + 1, q)
                if add:
                    np.add.at(x, ii, buf)
                else:
                    x[ii] = buf

    def comm(self, vecs):
        self._exchange(vecs, reverse=False, add=True)     # non-owners send partial sums, the owner adds
        self._exchange(vecs, reverse=True, add=False)     # the owner sends the total back
        return vecs

    def dot(self, xs, ys):
        return float(sum(np.dot(imp["weight"] * x, y) for imp, x, y in zip(self.imp, xs, ys)))

    # --- linear system ------------------------------------------------------------------------------
    def assemble(self, kind=KIND_POISSON, func=FUNC_LINEAR, fprm=None, mat=None):
        for s in self.spaces:
            s.assemble(kind, func, fprm, mat, sym=False, route=1)
        self.b = self.comm([s.rhs().copy() for s in self.spaces])   # par_sparse_assembler.f90:171-176
        return self.b

    def spmv(self, xs):
        return self.comm([s.spmv(x) for s, x in zip(self.spaces, xs)])

    def jacobi_setup(self):
        d = []
        for s in self.spaces:
            dinv = s.jacobi_setup()
            d.append(1.0 / dinv)                              # local (sub-assembled) diagonal
        self.dinv = [1.0 / v for v in self.comm(d)]
        return self.dinv

    def cg(self, prec=1, rtol=1e-6, atol=0.0, maxit=1000, stop=0):
        """cg.f90:101-238 on par_scalar_array_t; `stop` numbering as in fempar_oracle.cpp (cg.f90:279-407)."""
        b = self.b
        M = (lambda r: [d * v for d, v in zip(self.dinv, r)]) if prec == 1 else (lambda r: [v.copy() for v in r])
        nrm2 = lambda v: np.sqrt(self.dot(v, v))
        x = [np.zeros_like(v) for v in b]
        b_nrm_M = np.sqrt(self.dot(b, M(b))) if stop == 2 else 0.0
        r = [bv - av for bv, av in zip(b, self.spmv(x))]
        z = M(r)
        r_z = self.dot(r, z)
        r_nrm_M = np.sqrt(r_z)
        p = [v.copy() for v in z]
        tol = tol_extra = 0.0
        if stop in (4, 1):
            tol = rtol * nrm2(b) + atol
        elif stop in (7, 9):
            tol = tol_extra = rtol * nrm2(b) + atol
        elif stop == 0:
            tol = rtol * nrm2(r) + atol
        elif stop == 8:
            tol = tol_extra = rtol * nrm2(r) + atol
        elif stop == 6:
            tol, tol_extra = rtol * nrm2(b) + atol, rtol * nrm2(r) + atol
        elif stop == 2:
            tol = rtol * b_nrm_M + atol
        elif stop == 3:
            tol = rtol * r_nrm_M + atol
        it, conv, e, hist = 0, False, 0.0, []
        while True:
            it += 1
            Ap = self.spmv(p)
            App = self.dot(Ap, p)
            alpha = r_z / App if App != 0.0 else 0.0
            for xv, rv, pv, av in zip(x, r, p, Ap):
                xv += alpha * pv
                rv -= alpha * av
            delta = stop >= 4
            if it == 1 and delta:
                e = alpha * nrm2(p)
                if stop in (5, 8, 9):
                    tol = rtol * e + atol
            if stop in (0, 1):
                e = nrm2(r)
            elif delta:
                if it != 1:
                    e = alpha * nrm2(p)
            else:
                e = r_nrm_M
            hist.append(e)
            conv = e <= tol
            if stop >= 6 and conv:
                conv = nrm2(r) <= tol_extra
            if conv or it >= maxit:
                break
            z = M(r)
            beta = 1.0 / r_z
            r_z = self.dot(r, z)
            beta *= r_z
            r_nrm_M = np.sqrt(r_z)
            p = [zv + beta * pv for zv, pv in zip(z, p)]
        return dict(x=x, iters=it, converged=bool(conv), err=float(e), history=np.array(hist))

    def errors(self, xs, func=FUNC_LINEAR, fprm=None):
        h1 = l2 = 0.0
        for s, x in zip(self.spaces, xs):
            a, b = s.errors(x, func, fprm)
            h1 += a * a
            l2 += b * b
        return np.sqrt(h1), np.sqrt(l2)


def par_run_threads(spaces, kind=KIND_POISSON, func=FUNC_LINEAR, fprm=None, mat=None, cg_iters=25, steps=1, warmup=0):
    """The reference's MPI path with every rank a host thread (fo_par_run in fempar_oracle.cpp): sub-assembled parts,
    2-stage neighbour exchange through shared-memory mailboxes, weighted dots + all-reduce, `cg_iters` Jacobi-PCG iterations.
    Returns seconds per step/apply and the checksums |b|, |r|, sum diag(A)."""
    import ctypes as C
    L = spaces[0].L
    hs = (C.c_void_p * len(spaces))(*[s.h for s in spaces])
    a, pa = OracleSpace._prm(fprm)
    m, pm = OracleSpace._prm(mat)
    out = np.zeros(8)
    L.fo_par_run(hs, len(spaces), int(kind), int(func), C.c_void_p(pa), C.c_void_p(pm), int(cg_iters), int(steps), int(warmup),
                 C.c_void_p(out.ctypes.data))
    return dict(assembly_s=out[0], apply_s=out[1], pcg_s=out[2], jacobi_s=out[6], b_nrm2=out[3], r_nrm2=out[4], diag_sum=out[5])
