"""CPU tests of oracle/nonlinear_oracle.py (checker of the device residual / tangent path)."""
import numpy as np
import scipy.sparse as sp

from oracle.oracle import OracleSpace, KIND_POISSON, FUNC_SHARP_WAVE, FUNC_LINEAR
from oracle.nonlinear_oracle import residual_and_tangent

WAVE = [200.0, 0.7, -0.05, -0.05, -0.05]


def test_beta_zero_is_the_affine_operator():
    """beta = 0, f = 0: residual = A u - b and tangent = A of the linear oracle (Dirichlet data enter through u_h)"""
    N = (4, 3, 3)
    o = OracleSpace(N, order=2)
    o.assemble(KIND_POISSON, FUNC_LINEAR, route=1)
    A = sp.csr_matrix((o.val().copy(), (o.ja() - 1).copy(), o.irp().copy()), shape=(o.num_free, o.num_free))
    u = np.random.default_rng(0).standard_normal(o.num_free)
    r, T = residual_and_tangent(o, N, u, beta=0.0)
    assert np.max(np.abs(r - (A @ u - o.rhs()))) <= 1e-12 * np.max(np.abs(A @ u))
    assert np.array_equal(T.indptr, o.irp()) and np.array_equal(T.indices + 1, o.ja())
    assert np.max(np.abs(T.data - o.val())) <= 1e-12 * np.max(np.abs(o.val()))


def test_tangent_is_the_derivative_of_the_residual():
    N = (3, 3, 2)
    o = OracleSpace(N, order=2)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, WAVE, route=1)
    rng = np.random.default_rng(1)
    u, v = rng.standard_normal(o.num_free), rng.standard_normal(o.num_free)
    r, T = residual_and_tangent(o, N, u, beta=0.7)
    eps = 1e-6
    rp, _ = residual_and_tangent(o, N, u + eps * v, beta=0.7, tangent=False)
    rm, _ = residual_and_tangent(o, N, u - eps * v, beta=0.7, tangent=False)
    fd = (rp - rm) / (2 * eps)
    assert np.max(np.abs(T @ v - fd)) <= 1e-7 * np.max(np.abs(fd))
    assert np.max(np.abs((T - T.T).data)) > 1e-3 * np.max(np.abs(T.data))      # genuinely non-symmetric
