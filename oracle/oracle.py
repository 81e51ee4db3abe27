"""ctypes driver of the CPU oracle (oracle/fempar_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(fempar_b200/) must never import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libfempar_oracle.so")

FUNC_ZERO, FUNC_LINEAR, FUNC_SHARP_WAVE = 0, 1, 2
KIND_POISSON, KIND_VECTOR_LAPLACIAN, KIND_ELASTICITY = 0, 1, 2


def build(force=False):
    src = os.path.join(_HERE, "fempar_oracle.cpp")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        L.fo_create.restype = C.c_void_p
        L.fo_create.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint]
        for name in ("fo_lst_dofs_gids", "fo_lst_vefs", "fo_vef_bid", "fo_cells_gid", "fo_cells_part", "fo_vefs_gid",
                     "fo_irp", "fo_ja", "fo_val", "fo_rhs", "fo_fixed", "fo_dinv", "fo_weight",
                     "fo_imp_ptr"):
            if hasattr(L, name):
                getattr(L, name).restype = C.c_void_p
        L.fo_nnz.restype = C.c_longlong
        _lib = L
    return _lib


def _arr(ptr, n, dtype):
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=n)


class OracleSpace:
    """One part (or the serial problem) of the oracle: mesh + FE space + linear system."""

    def __init__(self, N, order=1, ncomp=1, parts=None, part_id=0, dims=3, dir_mask=0xFFFFFFFF):
        self.L = lib()
        self.dims = dims
        N = np.ascontiguousarray(np.asarray(N, dtype=np.int32))
        P = np.ascontiguousarray(np.asarray(parts if parts is not None else [1] * dims, dtype=np.int32))
        self._keep = (N, P)
        self.h = C.c_void_p(self.L.fo_create(dims, N.ctypes.data, P.ctypes.data, int(part_id), order, ncomp,
                                             C.c_uint(dir_mask)))
        self.order, self.ncomp = order, ncomp

    def set_threads(self, nt=0):
        """threads used by the oracle's SpMV (0 = all hardware threads)."""
        if nt <= 0:
            nt = self.L.fo_hw_threads()
        self.L.fo_set_threads(self.h, int(nt))
        return int(nt)

    def close(self):
        if self.h:
            self.L.fo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # sizes
    @property
    def num_cells(self): return self.L.fo_num_cells(self.h)
    @property
    def num_local_cells(self): return self.L.fo_num_local_cells(self.h)
    @property
    def ndofs_per_cell(self): return self.L.fo_ndofs_per_cell(self.h)
    @property
    def num_free(self): return self.L.fo_num_free(self.h)
    @property
    def num_free_ghost(self): return self.L.fo_num_free_ghost(self.h)
    @property
    def num_dirichlet(self): return self.L.fo_num_dirichlet(self.h)
    @property
    def nnz(self): return self.L.fo_nnz(self.h)

    @property
    def num_interior(self): return self.L.fo_num_interior(self.h)
    @property
    def num_interface(self): return self.L.fo_num_interface(self.h)

    def dof_import(self):
        """dof_import_t of this part: dict with list_rcv, rcv_ptrs (0-based offsets), unpack_idx (1-based), list_snd,
        snd_ptrs, pack_idx, weight."""
        sz = (C.c_int * 4)()
        self.L.fo_imp_sizes(self.h, sz)
        nr, ns, nu, npk = [int(v) for v in sz]
        g = lambda w, n: _arr(self.L.fo_imp_ptr(self.h, w), n, np.int32).copy()
        return dict(list_rcv=g(0, nr), rcv_ptrs=g(1, nr + 1), unpack_idx=g(2, nu), list_snd=g(3, ns), snd_ptrs=g(4, ns + 1),
                    pack_idx=g(5, npk), weight=_arr(self.L.fo_weight(self.h), self.num_free, np.float64).copy())

    def lst_dofs_gids(self):
        return _arr(self.L.fo_lst_dofs_gids(self.h), self.num_cells * self.ndofs_per_cell, np.int32).reshape(
            self.num_cells, self.ndofs_per_cell)

    def cells_gid(self): return _arr(self.L.fo_cells_gid(self.h), self.num_cells, np.int64)
    def cells_part(self): return _arr(self.L.fo_cells_part(self.h), self.num_cells, np.int32)

    @staticmethod
    def _prm(p):
        if p is None:
            return None, None
        a = np.ascontiguousarray(np.asarray(p, dtype=np.float64))
        return a, a.ctypes.data

    def assemble(self, kind=KIND_POISSON, func=FUNC_LINEAR, fprm=None, mat=None, sym=False, route=1):
        a, pa = self._prm(fprm)
        m, pm = self._prm(mat)
        self.L.fo_assemble(self.h, kind, func, C.c_void_p(pa), C.c_void_p(pm), int(sym), route)

    def build_pattern(self, sym=False):
        self.L.fo_build_pattern(self.h, int(sym))

    def irp(self): return _arr(self.L.fo_irp(self.h), self.num_free + 1, np.int64)
    def ja(self): return _arr(self.L.fo_ja(self.h), self.nnz, np.int32)
    def val(self): return _arr(self.L.fo_val(self.h), self.nnz, np.float64)
    def rhs(self): return _arr(self.L.fo_rhs(self.h), self.num_free, np.float64)
    def fixed(self): return _arr(self.L.fo_fixed(self.h), self.num_dirichlet, np.float64)

    def spmv(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(self.num_free)
        self.L.fo_spmv(self.h, C.c_void_p(x.ctypes.data), C.c_void_p(y.ctypes.data))
        return y

    def jacobi_setup(self):
        self.L.fo_jacobi_setup(self.h)
        return _arr(self.L.fo_dinv(self.h), self.num_free, np.float64)

    def cg(self, b=None, prec=1, rtol=1e-6, atol=0.0, maxit=1000, stop=0, history=False):
        b = np.ascontiguousarray(self.rhs() if b is None else b, dtype=np.float64)
        x = np.zeros(self.num_free)
        it, cv, err = C.c_int(0), C.c_int(0), C.c_double(0)
        hist = np.zeros(maxit) if history else None
        self.L.fo_cg(self.h, prec, C.c_double(rtol), C.c_double(atol), maxit, stop, C.c_void_p(b.ctypes.data),
                     C.c_void_p(x.ctypes.data), C.byref(it), C.byref(cv), C.byref(err),
                     C.c_void_p(hist.ctypes.data if history else None))
        out = dict(x=x, iters=it.value, converged=bool(cv.value), err=err.value)
        if history:
            out["history"] = hist[:it.value]
        return out

    def errors(self, x, func=FUNC_LINEAR, fprm=None):
        x = np.ascontiguousarray(x, dtype=np.float64)
        a, pa = self._prm(fprm)
        h1, l2 = C.c_double(0), C.c_double(0)
        self.L.fo_errors(self.h, C.c_void_p(x.ctypes.data), func, C.c_void_p(pa), C.byref(h1), C.byref(l2))
        return h1.value, l2.value

    NORM_NAMES = ("mean_norm", "l1_norm", "l2_norm", "lp_norm", "linfty_norm", "h1_seminorm", "h1_norm", "w1p_seminorm", "w1p_norm",
                  "w1infty_seminorm", "w1infty_norm")

    def errors_all(self, x, func=FUNC_LINEAR, fprm=None, exponent=None):
        """the eleven norms of par_test_poisson's check_solution for this part: (dict of finalised norms, raw accumulators[8])"""
        x = np.ascontiguousarray(x, dtype=np.float64)
        a, pa = self._prm(fprm)
        acc, out = np.zeros(8), np.zeros(11)
        self.L.fo_errors_all(self.h, C.c_void_p(x.ctypes.data), func, C.c_void_p(pa), C.c_double(exponent or 0.0),
                             C.c_void_p(acc.ctypes.data), C.c_void_p(out.ctypes.data))
        return dict(zip(self.NORM_NAMES, out)), acc

    @classmethod
    def finalize_norms(cls, accs, exponent=None):
        """l1_sum / l1_max of the per-part accumulators, then finalize_norm (error_norms.f90:174-221)"""
        accs = np.asarray(accs, dtype=np.float64).reshape(-1, 8)
        tot = accs.sum(axis=0)
        tot[4], tot[7] = accs[:, 4].max(), accs[:, 7].max()
        out = np.zeros(11)
        lib().fo_finalize_norms(C.c_void_p(tot.ctypes.data), C.c_double(exponent or 0.0), C.c_void_p(out.ctypes.data))
        return dict(zip(cls.NORM_NAMES, out))

    def element(self, cell, kind=KIND_POISSON, func=FUNC_LINEAR, fprm=None, mat=None):
        nd = self.ndofs_per_cell
        a, pa = self._prm(fprm)
        m, pm = self._prm(mat)
        elmat = np.zeros((nd, nd), order="F")
        elvec = np.zeros(nd)
        self.L.fo_element(self.h, cell, kind, func, C.c_void_p(pa), C.c_void_p(pm), C.c_void_p(elmat.ctypes.data),
                          C.c_void_p(elvec.ctypes.data))
        return elmat, elvec

    def ref_tables(self):
        nq = self.L.fo_nq(self.h)
        nn = self.ndofs_per_cell // self.ncomp
        qw = np.zeros(nq); phi = np.zeros((nq, nn)); dphi = np.zeros((nq, nn, 3))
        self.L.fo_ref_tables(self.h, C.c_void_p(qw.ctypes.data), C.c_void_p(phi.ctypes.data), C.c_void_p(dphi.ctypes.data))
        return qw, phi, dphi
