"""CPU oracle for host-numbered spaces on UNSTRUCTURED / distorted conforming hex meshes.

TEST INFRASTRUCTURE ONLY (same rule as oracle.py): imported by tests/ only, never by the product.

numpy restatement of the reference's cell integration with a real (non-affine) cell map:
  Jacobian of the trilinear map, inverse, |det|   Lib/Integration/CellMap/sbm_cell_map.i90:272-296
  physical gradients = J^-1 * reference gradients  Lib/Integration/CellIntegrators/sbm_cell_integrator.i90:448-561
  Poisson bilinear form  sum_q w|J| grad(phi_i).grad(phi_j)   Tests/test_poisson/.../poisson_cG_discrete_integration.f90
  elasticity bilinear form (lambda div div + 2 mu eps:eps)    Tests/par_test_... vector Laplacian / elasticity drivers
  strong Dirichlet columns moved to the RHS        sbm_fe_cell_iterator.i90:1818-1879
  COO sort + compress = scipy csr with summed duplicates and ascending columns (base_sparse_matrix.f90:4479-4628)
Tensor-product Gauss-Legendre rule with order+1 points per direction, equispaced Lagrange nodes, node/shape order
1 + i + j(p+1) + k(p+1)^2 -- the same reference FE fempar_oracle.cpp uses.

PINNING: on affine cells this must reproduce fempar_oracle.cpp (itself pinned on the reference's fixtures) entry by entry;
tests/test_oracle_hex.py checks exactly that, plus patch tests on distorted meshes.  There is no reference fixture with a
distorted mesh in /root/reference that can be run here (GiD meshes need the Fortran reader), so for non-affine cells parity is
"pinned through the affine case + exactness properties" only.
"""
import numpy as np
import scipy.sparse as sp


def lagrange_1d(p, x):
    """values L[q, i] and derivatives dL[q, i] of the equispaced Lagrange basis on [-1, 1] at the points x."""
    nodes = np.linspace(-1.0, 1.0, p + 1)
    x = np.asarray(x, dtype=np.float64)
    L = np.ones((x.size, p + 1))
    dL = np.zeros((x.size, p + 1))
    for i in range(p + 1):
        for j in range(p + 1):
            if j != i:
                L[:, i] *= (x - nodes[j]) / (nodes[i] - nodes[j])
        for k in range(p + 1):
            if k == i:
                continue
            t = np.full(x.size, 1.0 / (nodes[i] - nodes[k]))
            for j in range(p + 1):
                if j != i and j != k:
                    t *= (x - nodes[j]) / (nodes[i] - nodes[j])
            dL[:, i] += t
    return L, dL


def _tensor(Bx, By, Bz):
    """T[q, a] = Bx[q0, a0] By[q1, a1] Bz[q2, a2], x index fastest in both q and a."""
    t = np.einsum("xi,yj,zk->zyxkji", Bx, By, Bz)
    n = Bx.shape[0] * By.shape[0] * Bz.shape[0]
    return t.reshape(n, -1)


def reference_tables(p):
    xq, wq = np.polynomial.legendre.leggauss(p + 1)
    B, Cd = lagrange_1d(p, xq)
    G, Gd = lagrange_1d(1, xq)
    qw = np.einsum("x,y,z->zyx", wq, wq, wq).reshape(-1)
    phi = _tensor(B, B, B)
    dphi = np.stack([_tensor(Cd, B, B), _tensor(B, Cd, B), _tensor(B, B, Cd)], axis=-1)      # [q, a, 3]
    gphi = _tensor(G, G, G)
    gdphi = np.stack([_tensor(Gd, G, G), _tensor(G, Gd, G), _tensor(G, G, Gd)], axis=-1)     # [q, 8, 3]
    return qw, phi, dphi, gphi, gdphi


def cell_geometry(p, X):
    """X [ncell, 8, 3] -> physical q-points xq [c, q, 3], w|J| [c, q], physical gradients g [c, q, a, 3]."""
    qw, phi, dphi, gphi, gdphi = reference_tables(p)
    J = np.einsum("qai,caj->cqij", gdphi, X)                 # J[i, j] = d x_j / d xi_i
    Jinv = np.linalg.inv(J)                                   # Jinv[j, i] = d xi_i / d x_j
    det = np.abs(np.linalg.det(J))
    g = np.einsum("cqji,qai->cqaj", Jinv, dphi)
    xq = np.einsum("qa,cak->cqk", gphi, X)
    return xq, det * qw[None, :], g, phi


def element_matrices(p, X, ncomp=1, lam=0.0, mu=1.0, vector_laplacian=False):
    """element matrices [c, nd, nd]; ncomp = 1: Poisson; ncomp = 3: linear elasticity, or the vector Laplacian (the scalar
    matrix on the three diagonal blocks, explicit zeros elsewhere: vector_poisson_discrete_integration.f90:62-140);
    component-major DOF blocks."""
    xq, wd, g, phi = cell_geometry(p, X)
    if ncomp == 1:
        return np.einsum("cq,cqak,cqbk->cab", wd, g, g)
    nn = g.shape[2]
    K = np.zeros((X.shape[0], 3 * nn, 3 * nn))
    dot = np.einsum("cq,cqak,cqbk->cab", wd, g, g)
    if vector_laplacian:
        for ci in range(3):
            K[:, ci * nn:(ci + 1) * nn, ci * nn:(ci + 1) * nn] = dot
        return K
    for ci in range(3):
        for cj in range(3):
            blk = lam * np.einsum("cq,cqa,cqb->cab", wd, g[..., ci], g[..., cj]) + \
                  mu * np.einsum("cq,cqa,cqb->cab", wd, g[..., cj], g[..., ci])
            if ci == cj:
                blk = blk + mu * dot
            K[:, ci * nn:(ci + 1) * nn, cj * nn:(cj + 1) * nn] = blk
    return K


def node_coordinates(p, X):
    """physical positions [c, node, 3] of the Lagrange nodes under the trilinear map (nodal interpolation points)."""
    G, _ = lagrange_1d(1, np.linspace(-1.0, 1.0, p + 1))
    return np.einsum("na,cak->cnk", _tensor(G, G, G), X)


def assemble(p, ncomp, gids, X, n_free, fixed=None, source=None, lam=0.0, mu=1.0, sym=False, vector_laplacian=False):
    """CSR matrix (ascending columns, duplicates summed, explicit zeros kept) and RHS of the local cells (gids [c, nd], X [c, 8, 3])."""
    K = element_matrices(p, X, ncomp, lam, mu, vector_laplacian)
    nc, nd = gids.shape
    rhs = np.zeros(n_free)
    if source is not None:
        xq, wd, g, phi = cell_geometry(p, X)
        fv = source(xq)                                                   # [c, q]
        elvec = np.einsum("cq,qa->ca", wd * fv, phi)
    else:
        elvec = np.zeros((nc, nd))
    gi = np.repeat(gids[:, :, None], nd, axis=2)
    gj = np.repeat(gids[:, None, :], nd, axis=1)
    if fixed is not None and fixed.size:
        gfix = np.where(gids < 0, fixed[np.clip(-gids - 1, 0, fixed.size - 1)], 0.0)
        elvec = elvec - np.einsum("cab,cb->ca", K, gfix)
    rows_ok = (gids > 0) & (gids <= n_free)
    np.add.at(rhs, gids[rows_ok] - 1, elvec[rows_ok])
    keep = (gi > 0) & (gi <= n_free) & (gj > 0) & (gj <= n_free)
    if sym:
        keep &= gj >= gi
    A = sp.coo_matrix((K[keep], (gi[keep] - 1, gj[keep] - 1)), shape=(n_free, n_free)).tocsr()
    A.sum_duplicates()
    A.sort_indices()
    return A, rhs


def error_norms(p, gids, X, x, fixed, u, grad_u):
    """(H1 seminorm, L2 norm) of u - u_h over the given cells (tutorial_01_..._modules.f90:611-661)."""
    xq, wd, g, phi = cell_geometry(p, X)
    vals = np.where(gids > 0, x[np.clip(gids - 1, 0, max(x.size - 1, 0))], 0.0)
    if fixed is not None and fixed.size:
        vals = np.where(gids < 0, fixed[np.clip(-gids - 1, 0, fixed.size - 1)], vals)
    uh = np.einsum("ca,qa->cq", vals, phi)
    gh = np.einsum("ca,cqak->cqk", vals, g)
    e0 = np.sum(wd * (u(xq) - uh) ** 2)
    e1 = np.sum(wd * np.sum((grad_u(xq) - gh) ** 2, axis=-1))
    return np.sqrt(e1), np.sqrt(e0)


def structured_cell_coords(N, warp=None):
    """vertex coordinates [N^3, 8, 3] of the uniform unit-cube mesh (cells x fastest); `warp` maps points [.., 3] -> [.., 3]."""
    i = np.arange(N ** 3)
    ijk = np.stack([i % N, (i // N) % N, i // (N * N)], axis=1)
    a = np.arange(8)
    off = np.stack([a & 1, (a >> 1) & 1, (a >> 2) & 1], axis=1)
    X = (ijk[:, None, :] + off[None, :, :]) / float(N)
    return warp(X) if warp is not None else X


def y_junction_mesh(n=2, nz=2):
    """A genuinely unstructured conforming hex mesh: three quad blocks meeting at a valence-3 line (a 'Y' junction, the
    building block of O-grids), each block n x n cells, extruded into nz layers.  Returns X [3 n^2 nz, 8, 3]."""
    ang = lambda d: np.array([np.cos(np.deg2rad(d)), np.sin(np.deg2rad(d))])
    C0 = np.zeros(2)
    P = [ang(90), ang(210), ang(330)]
    M = [ang(150), ang(270), ang(30)]                       # M[0] between P0 and P1, M[1] between P1 and P2, M[2] between P2, P0
    blocks = [(C0, M[2], P[0], M[0]), (C0, M[0], P[1], M[1]), (C0, M[1], P[2], M[2])]   # counter-clockwise corners
    cells = []
    for (A, B, Cc, D) in blocks:
        bil = lambda s, t: A * (1 - s) * (1 - t) + B * s * (1 - t) + D * (1 - s) * t + Cc * s * t
        for k in range(nz):
            for j in range(n):
                for i in range(n):
                    v = np.zeros((8, 3))
                    for a in range(8):
                        s, t, z = (i + (a & 1)) / n, (j + ((a >> 1) & 1)) / n, (k + ((a >> 2) & 1)) / nz
                        v[a, :2] = bil(s, t)
                        v[a, 2] = z
                    cells.append(v)
    return np.array(cells)


def number_by_coordinates(p, X, ncomp=1, decimals=9):
    """Conforming H1 numbering of any trilinear hex mesh: Lagrange nodes that coincide physically are one DOF, ids handed out
    by first touch in cell order (cell-local node order); nodes of faces owned by a single cell are strong Dirichlet.
    Returns gids [c, ncomp * (p+1)^3] (free > 0, Dirichlet < 0), n_free, n_dirichlet.  (The device accepts ANY numbering the
    host supplies; this one only has to be conforming.)"""
    xn = node_coordinates(p, X)
    nc, nn, _ = xn.shape
    key = np.round(xn.reshape(-1, 3), decimals)
    _, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    order = np.argsort(np.argsort(first))                    # unique node -> rank by first touch
    node = order[inv].reshape(nc, nn)
    n1 = p + 1
    loc = np.arange(nn)
    ijk = np.stack([loc % n1, (loc // n1) % n1, loc // (n1 * n1)], axis=1)
    faces = {}
    face_nodes = []
    for d in range(3):
        for side in (0, p):
            face_nodes.append(loc[ijk[:, d] == side])
    for c in range(nc):
        for f in face_nodes:
            k = tuple(sorted(node[c, f]))
            faces[k] = faces.get(k, 0) + 1
    is_dir = np.zeros(order.size, dtype=bool)
    for k, cnt in faces.items():
        if cnt == 1:
            is_dir[list(k)] = True
    gids = np.zeros((nc, ncomp * nn), dtype=np.int32)
    free_id = np.cumsum(~is_dir) * (~is_dir)                 # 1-based among free nodes (first-touch order)
    dir_id = np.cumsum(is_dir) * is_dir
    nf, ndir = int((~is_dir).sum()), int(is_dir.sum())
    for cc in range(ncomp):
        g = np.where(is_dir[node], -(dir_id[node] + cc * ndir), free_id[node] + cc * nf)
        gids[:, cc * nn:(cc + 1) * nn] = g
    return gids, ncomp * nf, ncomp * ndir
