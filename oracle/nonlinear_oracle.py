"""TEST INFRASTRUCTURE ONLY (checker; never imported by the product).

numpy restatement of what fe_operator_t does for a non-linear operator at an evaluation point
(/root/reference/Sources/Lib/Integration/FEOperator/sbm_fe_operator.i90:568-639: compute_internal_residual ->
discrete_integration%integrate_residual, compute_tangent -> integrate_tangent), with the cell loop of the reference's own
integrate_residual (Sources/Tests/Serial/test_transient_poisson/modules/test_transient_poisson_discrete_integration.f90:
248-319: u_h%update(fe, fe_function), gradients at the quadrature points, elvec(i) += factor * (grad phi_i . grad u_h - f phi_i),
fe%assembly) extended by the conductivity k(u) = 1 + beta u^2 and its Gateaux derivative for the tangent.
Parity status: beta = 0 is pinned on the C++ oracle (residual = A u - b, tangent = A, tests/test_oracle_nonlinear.py); for
beta /= 0 the reference holds no integrand of this kind, the tangent is pinned on central differences of the residual.
Geometry: cells of the structured unit-cube mesh of OracleSpace (affine, J = diag(h/2) for the reference cell [-1,1]^3).
"""
import numpy as np
import scipy.sparse as sp


def residual_and_tangent(o, N, u_free, beta=0.0, f_q=None, tangent=True):
    """o: serial OracleSpace (assembled once so that fixed() holds the Dirichlet values); N: cells per direction;
    u_free: free DOF values of the evaluation point; f_q: source samples [cell][q] (None: f = 0).
    Returns (residual[n_free], tangent csr[n_free, n_free] or None)."""
    gids = o.lst_dofs_gids()
    fixed = o.fixed()
    qw, phi, dphi = o.ref_tables()
    h = 1.0 / np.asarray(N, dtype=np.float64)
    J = h / 2.0
    w = qw * np.prod(J)
    G = dphi / J[None, None, :]                      # physical gradients of the shape functions [q, i, 3]
    n = o.num_free
    r = np.zeros(n)
    rows, cols, vals = [], [], []
    for c in range(o.num_local_cells):
        g = gids[c]
        ul = np.where(g > 0, u_free[np.maximum(g, 1) - 1], fixed[np.maximum(-g, 1) - 1])
        uq = phi @ ul
        gu = np.einsum("qik,i->qk", G, ul)
        k, kp = 1.0 + beta * uq * uq, 2.0 * beta * uq
        fq = f_q[c] if f_q is not None else np.zeros_like(uq)
        rl = np.einsum("q,qik,qk->i", w * k, G, gu) - phi.T @ (w * fq)
        free = (g > 0) & (g <= n)
        np.add.at(r, g[free] - 1, rl[free])
        if tangent:
            T = np.einsum("q,qik,qjk->ij", w * k, G, G) + np.einsum("q,qik,qk,qj->ij", w * kp, G, gu, phi)
            ii, jj = np.nonzero(np.outer(free, free))
            rows.append(g[ii] - 1); cols.append(g[jj] - 1); vals.append(T[ii, jj])
    Tm = None
    if tangent:
        Tm = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(n, n)).tocsr()
        Tm.sum_duplicates(); Tm.sort_indices()
    return r, Tm
