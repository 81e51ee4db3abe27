"""Pins oracle/hex_oracle.py (distorted / unstructured hex restatement): on affine cells it must reproduce the
pinned C++ oracle entry by entry; on distorted cells the isoparametric patch test must hold."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse.linalg as spla

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import hex_oracle as ho                                            # noqa: E402
from oracle.oracle import OracleSpace, KIND_POISSON, KIND_ELASTICITY, FUNC_LINEAR, FUNC_SHARP_WAVE   # noqa: E402


def warp(X):
    x, y, z = X[..., 0], X[..., 1], X[..., 2]
    return np.stack([x + 0.06 * np.sin(np.pi * x) * np.sin(2 * np.pi * y) * (1 + z),
                     y + 0.05 * np.sin(np.pi * y) * np.sin(np.pi * z) * (1 - 0.5 * x),
                     z + 0.04 * np.sin(np.pi * z) * np.cos(np.pi * x) * y], axis=-1)


@pytest.mark.parametrize("p", [1, 2, 3])
def test_reference_tables_match_pinned_oracle(p):
    o = OracleSpace([2] * 3, order=p)
    qw, phi, dphi = o.ref_tables()
    hq, hphi, hdphi, _, _ = ho.reference_tables(p)
    # the C++ oracle tabulates on its own reference interval: compare through interval-independent quantities
    s = hq.sum() / qw.sum()
    assert np.allclose(hq, qw * s, rtol=0, atol=1e-14 * s)
    assert np.allclose(hphi, phi, rtol=0, atol=1e-13)
    L = s ** (1.0 / 3.0)                                                       # ratio of the interval lengths
    assert np.allclose(hdphi * L, dphi, rtol=0, atol=1e-12 * max(1.0, np.abs(dphi).max()))


@pytest.mark.parametrize("p,sym", [(1, False), (2, False), (2, True), (3, False)])
def test_affine_poisson_matches_pinned_oracle(p, sym):
    N = 3
    o = OracleSpace([N] * 3, order=p)
    o.assemble(KIND_POISSON, FUNC_SHARP_WAVE, [200, .7, -.05, -.05, -.05], None, sym=sym, route=1)
    gids = o.lst_dofs_gids().copy()
    alpha, r, c = 200.0, 0.7, -0.05

    def f(x):                                                                  # analytical_functions of tutorial_01 (source)
        d2 = (x - c) ** 2
        t1 = d2.sum(-1); s = np.sqrt(t1); t2 = s - r; w = alpha ** 2 * t2 ** 2 + 1.0
        res = 0.0
        for k in range(3):
            res = res - (alpha / (s * w) - alpha * d2[..., k] / (t1 * s * w) - 2 * alpha ** 3 * d2[..., k] * t2 / (t1 * w * w))
        return res
    A, rhs = ho.assemble(p, 1, gids, ho.structured_cell_coords(N), o.num_free, o.fixed().copy(), source=f, sym=sym)
    assert np.array_equal(A.indptr, o.irp()) and np.array_equal(A.indices + 1, o.ja())
    scale = np.abs(o.val()).max()
    assert np.abs(A.data - o.val()).max() <= 1e-12 * scale
    assert np.abs(rhs - o.rhs()).max() <= 1e-11 * max(1.0, np.abs(o.rhs()).max())


def test_affine_elasticity_matches_pinned_oracle():
    N, lam, mu = 3, 1.3, 0.7
    o = OracleSpace([N] * 3, order=1, ncomp=3)
    o.assemble(KIND_ELASTICITY, FUNC_LINEAR, None, [lam, mu], sym=False, route=1)
    A, rhs = ho.assemble(1, 3, o.lst_dofs_gids().copy(), ho.structured_cell_coords(N), o.num_free, o.fixed().copy(),
                         lam=lam, mu=mu)
    assert np.array_equal(A.indptr, o.irp()) and np.array_equal(A.indices + 1, o.ja())
    assert np.abs(A.data - o.val()).max() <= 1e-12 * np.abs(o.val()).max()
    assert np.abs(rhs - o.rhs()).max() <= 1e-12 * max(1.0, np.abs(o.rhs()).max())


@pytest.mark.parametrize("p", [1, 2])
def test_distorted_patch_test(p):
    """isoparametric elements reproduce u = x + y + z exactly on any trilinear hex mesh."""
    N = 4
    o = OracleSpace([N] * 3, order=p)
    gids = o.lst_dofs_gids().copy()
    X = ho.structured_cell_coords(N, warp)
    xn = ho.node_coordinates(p, X)
    fixed = np.zeros(o.num_dirichlet)
    u_nodes = xn.sum(-1)
    fixed[-gids[gids < 0] - 1] = u_nodes[gids < 0]
    A, rhs = ho.assemble(p, 1, gids, X, o.num_free, fixed)
    x = spla.spsolve(A.tocsc(), rhs)
    exact = np.zeros(o.num_free)
    exact[gids[gids > 0] - 1] = u_nodes[gids > 0]
    assert np.abs(x - exact).max() < 1e-11
    h1, l2 = ho.error_norms(p, gids, X, x, fixed, lambda q: q.sum(-1), lambda q: np.ones_like(q))
    assert h1 < 1e-10 and l2 < 1e-11


@pytest.mark.parametrize("p", [1, 2])
def test_y_junction_mesh_patch_test(p):
    """unstructured topology (a line where three blocks meet): conforming numbering + exact reproduction of u = x+y+z."""
    X = ho.y_junction_mesh(3, 2)
    gids, nf, ndir = ho.number_by_coordinates(p, X)
    assert nf > 0 and ndir > 0 and gids.max() == nf and gids.min() == -ndir
    vol = ho.cell_geometry(p, X)[1].sum()
    assert abs(vol - 1.5 * np.sqrt(3.0)) < 1e-12            # regular hexagon of circumradius 1, height 1
    xn = ho.node_coordinates(p, X).sum(-1)
    fixed = np.zeros(ndir)
    fixed[-gids[gids < 0] - 1] = xn[gids < 0]
    A, rhs = ho.assemble(p, 1, gids, X, nf, fixed)
    assert abs(A - A.T).max() < 1e-13
    x = spla.spsolve(A.tocsc(), rhs)
    exact = np.zeros(nf)
    exact[gids[gids > 0] - 1] = xn[gids > 0]
    assert np.abs(x - exact).max() < 1e-11
    valence = np.bincount(gids[:, 0][gids[:, 0] > 0])      # local node 0 of every block cell (i,j)=(0,0) is the junction line
    assert valence.max() == 3


def test_affine_vector_laplacian_matches_pinned_oracle():
    from oracle.oracle import KIND_VECTOR_LAPLACIAN
    N = 3
    o = OracleSpace([N] * 3, order=2, ncomp=3)
    o.assemble(KIND_VECTOR_LAPLACIAN, FUNC_LINEAR, None, None, sym=False, route=1)
    A, rhs = ho.assemble(2, 3, o.lst_dofs_gids().copy(), ho.structured_cell_coords(N), o.num_free, o.fixed().copy(),
                         vector_laplacian=True)
    assert np.array_equal(A.indptr, o.irp()) and np.array_equal(A.indices + 1, o.ja())     # explicit zeros kept on both sides
    assert np.abs(A.data - o.val()).max() <= 1e-12 * np.abs(o.val()).max()
    assert np.abs(rhs - o.rhs()).max() <= 1e-12 * max(1.0, np.abs(o.rhs()).max())
