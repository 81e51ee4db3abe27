"""TEST INFRASTRUCTURE ONLY (checker, never shipped or measured as the product).

CPU restatement (numpy) of the reference's other Krylov solvers that share the CG path's primitives
(SURVEY.md 8f row 3), each following /root/reference/Sources/Lib/LinearAlgebra/IterativeLinearSolver/:
  richardson.f90:101-205, icg.f90:106-240 (+ the same convergence bookkeeping as cg.f90:279-407),
  rgmres.f90:120-356, fgmres.f90:120-356, lgmres.f90:152-420,
  iterative_solver_utils.f90:46-106 (modified_gs_reorthonorm), :128-196 (iterative_gs_reorthonorm),
  :202-235 (apply_givens_rotation).
Parity status: the reference holds no golden vectors for these solvers (its tests only check that the solution error is below
a tolerance), so the restatement is pinned on an INDEPENDENT implementation instead (tests/test_oracle_krylov.py): scipy 1.18
GMRES (per-iteration residual norms and iterates of R/F/LGMRES, identity and Jacobi preconditioning, with restarts) and
MINRES (iterates after k iterations, with and without M) to 1e-8 .. 1e-9, LFOM against a dense Arnoldi projection and
Richardson against the written-out recurrence.  What stays unpinned against the reference itself: its exact choice of
error estimate per stopping criterion (cited line by line).

A, M are callables y = A(x), z = M(r) on 1-D float64 arrays; dot defaults to numpy's.
"""
import numpy as np

MGSRO, ICGSRO = 0, 1          # orthonorm_strat_mgsro / orthonorm_strat_icgsro (iterative_solver_parameters.f90)
# stopping criteria numbering shared with fempar_oracle.cpp / include/fempar_b200.h
RES_RES, RES_RHS, RES_NRMGIVEN_RHS_NRMGIVEN, RES_NRMGIVEN_RES_NRMGIVEN = 0, 1, 2, 3


def _nrm2(v):
    return float(np.sqrt(np.dot(v, v)))


def apply_givens_rotation(x, y):
    """iterative_solver_utils.f90:202-235 (SPARSKIT): returns (x', y', c, s) with (x, y) -> (|r|, 0)."""
    if x == 0.0 and y == 0.0:
        c, s = 1.0, 0.0
    elif abs(y) > abs(x):
        t = x / y
        x = np.sqrt(1.0 + t * t)
        s = np.copysign(1.0 / x, y)
        c = t * s
    elif abs(y) <= abs(x):
        t = y / x
        y = np.sqrt(1.0 + t * t)
        c = np.copysign(1.0 / y, x)
        s = t * c
    else:
        x = y = 0.0
        c, s = 1.0, 0.0
    return abs(x * y), 0.0, c, s


def modified_gs_reorthonorm(m, V, hh):
    """iterative_solver_utils.f90:46-106.  V[m-1] is orthonormalised against V[0..m-2]; hh[0..m-1] filled; returns ierrc."""
    reorth = np.float64(np.float32(0.98))      # `reorth = 0.98` is a default-kind (single) literal in the reference
    q = V[m - 1]
    norm = float(np.dot(q, q))
    if norm <= 0.0:
        return -1
    if not (norm > 0.0 and 1.0 / norm > 0.0):
        return -2
    thr = norm * reorth
    for i in range(m - 1):
        fct = float(np.dot(q, V[i]))
        hh[i] = fct
        q -= fct * V[i]
        if fct * fct > thr:
            fct = float(np.dot(q, V[i]))
            hh[i] += fct
            q -= fct * V[i]
        norm -= hh[i] * hh[i]
        if norm < 0.0:
            norm = 0.0
        thr = norm * reorth
    hh[m - 1] = _nrm2(q)
    if hh[m - 1] <= 0.0:
        return -3
    q *= 1.0 / hh[m - 1]
    return 0


def iterative_gs_reorthonorm(k, Q, s):
    """iterative_solver_utils.f90:128-196 (iterated classical Gram-Schmidt, multidot + multiaxpy)."""
    alpha = 0.5
    s[:k] = 0.0
    q = Q[k - 1]
    delta_prev = _nrm2(q)
    if delta_prev <= 0.0:
        return -1
    if not (delta_prev > 0.0 and 1.0 / delta_prev > 0.0):
        return -2
    while True:
        p = np.array([float(np.dot(Q[i], q)) for i in range(k - 1)])
        for i in range(k - 1):
            q -= p[i] * Q[i]
        s[:k - 1] += p
        delta = _nrm2(q)
        if delta > delta_prev * alpha or delta == 0.0:
            break
        delta_prev = delta
    s[k - 1] = delta
    if s[k - 1] <= 0.0:
        return -3
    q *= 1.0 / s[k - 1]
    return 0


def _ortho(strat, m, V, hh):
    return modified_gs_reorthonorm(m, V, hh) if strat == MGSRO else iterative_gs_reorthonorm(m, V, hh)


def richardson(A, M, b, rtol=1e-6, atol=0.0, maxit=1000, stop=RES_RES, relaxation=1.0):
    """richardson.f90:101-205 (supports res_res, res_rhs)."""
    x = np.zeros_like(b)
    r = b - A(x)
    res = _nrm2(r)
    tol = rtol * (_nrm2(b) if stop == RES_RHS else res) + atol
    err, hist, it = res, [], 0
    conv = err <= tol
    while not conv and it < maxit:
        x = x + relaxation * M(r)
        it += 1
        r = b - A(x)
        err = _nrm2(r)
        hist.append(err)
        conv = err <= tol
    return dict(x=x, iters=it, converged=bool(conv), err=err, history=np.array(hist))


def icg(A, M, b, rtol=1e-6, atol=0.0, maxit=1000, stop=RES_RES):
    """icg.f90:106-240: CG with the Polak-Ribiere beta = (<r,z> - <r_old,z>)/<r_old,z_old>; convergence bookkeeping
    identical to cg.f90:279-407 (icg.f90:283-413), all ten criteria (numbering as in fempar_oracle.cpp)."""
    nrm2 = _nrm2
    x = np.zeros_like(b)
    b_nrm_M = np.sqrt(np.dot(b, M(b))) if stop == 2 else 0.0
    r = b - A(x)
    z = M(r)
    r_z = float(np.dot(r, z))
    r_nrm_M = np.sqrt(r_z)
    p = z.copy()
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
        Ap = A(p)
        App = float(np.dot(Ap, p))
        alpha = r_z / App if App != 0.0 else 0.0
        x = x + alpha * p
        r2 = r.copy()
        r = r - alpha * Ap
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
        r_z = float(np.dot(r, z))
        r2_z = float(np.dot(r2, z))
        beta = beta * (r_z - r2_z)
        r_nrm_M = np.sqrt(r_z)
        p = z + beta * p
    return dict(x=x, iters=it, converged=bool(conv), err=float(e), history=np.array(hist))


def _back_substitute(hh, g, kloc):
    """solve hh(1:kloc,1:kloc) y = g in place (rgmres.f90:330-337, the non-BLAS branch = DTRSV 'U','N','N')."""
    for j in range(kloc - 1, -1, -1):
        g[j] = g[j] / hh[j, j]
        for i in range(j - 1, -1, -1):
            g[i] -= hh[i, j] * g[j]


def _gmres(A, M, b, rtol, atol, maxit, stop, dkrymax, strat, variant):
    """Shared restatement of rgmres.f90:120-356 ('R': right preconditioning, x += M(V y)), fgmres.f90:120-356 ('F':
    flexible, the preconditioned basis Z is stored, x += Z y) and lgmres.f90:152-420 ('L': left preconditioning,
    Krylov space of M A, residual norm = ||M r||; additionally supports res_res / res_rhs by forming the true residual
    every iteration)."""
    n = b.size
    m = dkrymax
    x = np.zeros(n)
    V = [None] * (m + 1)
    Z = [None] * (m + 1)
    hh = np.zeros((m + 1, m))
    g = np.zeros(m + 1)
    cs = np.zeros((2, m))
    rhs_norm = _nrm2(b) if (stop == RES_NRMGIVEN_RHS_NRMGIVEN or (variant == 'L' and stop == RES_RHS)) else 0.0
    r = b - A(x)
    res_norm = _nrm2(r)
    res_2_norm = res_norm
    if variant == 'L':
        z = M(r)
        res_norm = _nrm2(z)
    if stop == RES_NRMGIVEN_RHS_NRMGIVEN:
        tol, err = rtol * rhs_norm + atol, res_norm
    elif stop == RES_NRMGIVEN_RES_NRMGIVEN:
        tol, err = rtol * res_norm + atol, res_norm
    elif stop == RES_RES:
        assert variant == 'L'
        tol, err = rtol * res_2_norm + atol, res_2_norm
    else:
        assert variant == 'L' and stop == RES_RHS
        tol, err = rtol * rhs_norm + atol, res_2_norm
    exit_loop = err <= tol
    it, hist, ierrc = 0, [], 0
    while not exit_loop and it < maxit:
        if it != 0:
            r = b - A(x)
            if variant == 'L':
                z = M(r)
                res_norm = _nrm2(z)
            else:
                res_norm = _nrm2(r)
        v0 = z if variant == 'L' else r
        V[0] = (v0 / res_norm) if res_norm != 0.0 else x.copy()   # bkryv%clone(x): x's values are not copied by clone in
        g[:] = 0.0                                                # the reference; res_norm == 0 never enters the loop
        g[0] = res_norm
        kloc = 0
        z_lg = None
        while not exit_loop and it < maxit and kloc < m:
            kloc += 1
            it += 1
            if variant == 'R':
                V[kloc] = A(M(V[kloc - 1]))
            elif variant == 'F':
                Z[kloc - 1] = M(V[kloc - 1])
                V[kloc] = A(Z[kloc - 1])
            else:
                V[kloc] = M(A(V[kloc - 1]))
            col = np.zeros(kloc + 1)
            ierrc = _ortho(strat, kloc + 1, V, col)
            hh[:kloc + 1, kloc - 1] = col
            if ierrc < 0:
                exit_loop = True
                break
            for j in range(kloc - 1):
                a = hh[j, kloc - 1]
                c, s = cs[0, j], cs[1, j]
                hh[j, kloc - 1] = c * a + s * hh[j + 1, kloc - 1]
                hh[j + 1, kloc - 1] = c * hh[j + 1, kloc - 1] - s * a
            hh[kloc - 1, kloc - 1], hh[kloc, kloc - 1], c, s = apply_givens_rotation(hh[kloc - 1, kloc - 1], hh[kloc, kloc - 1])
            cs[0, kloc - 1], cs[1, kloc - 1] = c, s
            a = -s * g[kloc - 1]
            g[kloc - 1] = c * g[kloc - 1]
            g[kloc] = a
            res_norm = abs(a)
            if variant == 'L' and stop in (RES_RES, RES_RHS):
                ka = kloc
                ga = g[:ka].copy()
                while ka > 0 and hh[ka - 1, ka - 1] == 0.0:
                    ka -= 1
                if ka <= 0:
                    return dict(x=x, iters=it, converged=False, err=err, history=np.array(hist))
                _back_substitute(hh, ga, ka)
                z_lg = x.copy()
                for i in range(ka):
                    z_lg += ga[i] * V[i]
                res_2_norm = _nrm2(b - A(z_lg))
                err = res_2_norm
            else:
                err = res_norm
            hist.append(err)
            exit_loop = err <= tol
        if ierrc == -2:
            break
        if variant == 'L' and stop in (RES_RES, RES_RHS):
            if z_lg is not None:
                x = z_lg.copy()
        elif kloc > 0:
            while kloc > 0 and hh[kloc - 1, kloc - 1] == 0.0:
                kloc -= 1
            if kloc <= 0:
                break
            _back_substitute(hh, g, kloc)
            upd = np.zeros(n)
            basis = Z if variant == 'F' else V
            for i in range(kloc):
                upd += g[i] * basis[i]
            x = x + (M(upd) if variant == 'R' else upd)
        exit_loop = err <= tol
    return dict(x=x, iters=it, converged=bool(err <= tol), err=float(err), history=np.array(hist))


def rgmres(A, M, b, rtol=1e-6, atol=0.0, maxit=1000, stop=RES_NRMGIVEN_RES_NRMGIVEN, dkrymax=30, strat=ICGSRO):
    return _gmres(A, M, b, rtol, atol, maxit, stop, dkrymax, strat, 'R')


def fgmres(A, M, b, rtol=1e-6, atol=0.0, maxit=1000, stop=RES_NRMGIVEN_RES_NRMGIVEN, dkrymax=30, strat=ICGSRO):
    return _gmres(A, M, b, rtol, atol, maxit, stop, dkrymax, strat, 'F')


def lgmres(A, M, b, rtol=1e-6, atol=0.0, maxit=1000, stop=RES_NRMGIVEN_RES_NRMGIVEN, dkrymax=30, strat=ICGSRO):
    return _gmres(A, M, b, rtol, atol, maxit, stop, dkrymax, strat, 'L')


def lfom(A, M, b, rtol=1e-6, atol=0.0, maxit=1000, stop=RES_RES, dkrymax=30, strat=ICGSRO):
    """lfom.f90:124-330: left-preconditioned full orthogonalisation method; every iteration solves H_k y = beta e_1 by LU with
    partial pivoting (dgetrf/dgetrs: LAPACK, third party, version unpinned; numpy.linalg.solve is the same algorithm) and
    forms the true residual.  Supports res_res, res_rhs.  NOTE restated as written: g(1) is the INITIAL preconditioned
    residual norm also after a restart (lfom.f90:247), although the restarted basis is normalised with the current one."""
    n = b.size
    m = dkrymax
    x = np.zeros(n)
    V = [None] * (m + 1)
    rhs_norm = _nrm2(b) if stop == RES_RHS else 0.0
    r = b - A(x)
    res_2_norm = _nrm2(r)
    z = M(r)
    res_norm = _nrm2(z)
    res_norm_initial = res_norm
    tol = rtol * (res_2_norm if stop == RES_RES else rhs_norm) + atol
    err = res_2_norm
    exit_loop = err <= tol
    it, hist, ierrc = 0, [], 0
    zsol = x.copy()
    while not exit_loop and it < maxit:
        hh = np.zeros((m + 1, m))
        if it != 0:
            r = b - A(x)
            z = M(r)
            res_norm = _nrm2(z)
        V[0] = z / res_norm if res_norm != 0.0 else x.copy()
        kloc = 0
        while not exit_loop and it < maxit and kloc < m:
            kloc += 1
            it += 1
            V[kloc] = M(A(V[kloc - 1]))
            col = np.zeros(kloc + 1)
            ierrc = _ortho(strat, kloc + 1, V, col)
            hh[:kloc + 1, kloc - 1] = col
            if ierrc < 0:
                exit_loop = True
                break
            g = np.zeros(kloc)
            g[0] = res_norm_initial
            try:
                y = np.linalg.solve(hh[:kloc, :kloc], g)
            except np.linalg.LinAlgError:
                return dict(x=x, iters=it, converged=False, err=float(err), history=np.array(hist))
            zsol = x.copy()
            for i in range(kloc):
                zsol += y[i] * V[i]
            res_2_norm = _nrm2(b - A(zsol))
            err = res_2_norm
            hist.append(err)
            exit_loop = err <= tol
        if ierrc == -2:
            break
        x = zsol.copy()
        exit_loop = err <= tol
    return dict(x=x, iters=it, converged=bool(err <= tol), err=float(err), history=np.array(hist))


def minres(A, M, b, rtol=1e-6, atol=0.0, maxit=1000):
    """minres.f90:361-740 (Paige & Saunders' MINRES as adapted in the reference; only res_res is accepted, and the test
    actually applied is ||r_k||_inv(M) estimate phibar <= Anorm*ynorm*rtol + atol).  Returns istop as well."""
    eps = np.finfo(np.float64).eps
    x = np.zeros_like(b)
    istop, it = 0, 0
    Anorm = Acond = rnorm = ynorm = 0.0
    hist = []
    r1 = b - A(x)
    v1 = M(r1)
    beta1 = float(np.dot(r1, v1))
    err, tol = 0.0, 0.0
    if beta1 < 0.0:
        istop = 8
    elif beta1 == 0.0:
        istop = 0
    else:
        beta1 = np.sqrt(beta1)
        oldb, beta, dbar, epsln = 0.0, beta1, 0.0, 0.0
        qrnorm = phibar = rhs1 = beta1
        rhs2 = tnorm2 = ynorm2 = 0.0
        cs, sn = -1.0, 0.0
        w = np.zeros_like(b)
        w2 = np.zeros_like(b)
        r2 = r1.copy()
        gmax = gmin = 0.0
        while True:
            it += 1
            s = 1.0 / beta
            v2 = s * v1
            y = A(v2)
            if it >= 2:
                y = y - (beta / oldb) * r1
            alpha = float(np.dot(v2, y))
            y = y - (alpha / beta) * r2
            r1 = r2.copy()
            r2 = y.copy()
            v1 = M(r2)
            oldb = beta
            beta = float(np.dot(r2, v1))
            if beta < 0.0:
                istop = 6
                break
            beta = np.sqrt(beta)
            tnorm2 = tnorm2 + alpha ** 2 + oldb ** 2 + beta ** 2
            if it == 1:
                if beta / beta1 <= 10.0 * eps:
                    istop = -1
                gmax = abs(alpha)
                gmin = gmax
            oldeps = epsln
            delta = cs * dbar + sn * alpha
            gbar = sn * dbar - cs * alpha
            epsln = sn * beta
            dbar = -cs * beta
            gamma = np.sqrt(gbar ** 2 + beta ** 2)
            cs = gbar / gamma
            sn = beta / gamma
            phi = cs * phibar
            phibar = sn * phibar
            denom = 1.0 / gamma
            w1 = w2.copy()
            w2 = w.copy()
            w = (v2 - oldeps * w1 - delta * w2) * denom
            x = x + phi * w
            gmax = max(gmax, gamma)
            gmin = min(gmin, gamma)
            z = rhs1 / gamma
            ynorm2 = z ** 2 + ynorm2
            rhs1 = rhs2 - delta * z
            rhs2 = -epsln * z
            Anorm = np.sqrt(tnorm2)
            ynorm = np.sqrt(ynorm2)
            epsx = Anorm * ynorm * eps
            epsr = Anorm * ynorm * rtol + atol
            tol, err = epsr, phibar
            qrnorm = phibar
            rnorm = qrnorm
            Acond = gmax / gmin
            if istop == 0:
                if it >= maxit:
                    istop = 5
                if Acond >= 0.1 / eps:
                    istop = 4
                if epsx >= beta1:
                    istop = 3
                if qrnorm <= epsx:
                    istop = 2
                if qrnorm <= epsr:
                    istop = 1
            hist.append(err)
            if istop != 0:
                break
    conv = istop in (0, -1, 1, 2)
    return dict(x=x, iters=it, converged=bool(conv), err=float(err), history=np.array(hist), istop=istop,
                Anorm=float(Anorm), Acond=float(Acond))
