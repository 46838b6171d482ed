"""Pins oracle/p2_oracle.py (P2 function space, blended on-the-fly element matrices, IcosahedralShellMap) to OUTPUTS OF THE REFERENCE
ITSELF (tests/golden/ref_p2.npz): P2ElementwiseOperator< p2_div_k_grad_blending_q4 > and < p2_epsilonvar_i_j_blending_q2 >::apply on the
thick spherical shell of P2P1ElementwiseStokesOperatorTest.cpp (meshSphericalShell(3,2,0.5,1.0)) with the shell map, the affine case on
cube_6el, and evalF / evalDF samples of every macro-cell's map (skew tets included)."""
import numpy as np
import pytest

from hyteg_b200 import meshes
from oracle import p2_oracle as p2
import ref_cases as rc

TOL = 1e-12


@pytest.mark.parametrize("nt,nr", [(3, 2), (2, 3)])
def test_shell_map(nt, nr):
    from scipy.spatial import cKDTree
    g = rc.gold_p2()["shellmap_%d_%d" % (nt, nr)]
    v, c = meshes.spherical_shell(nt, nr, 0.5, 1.0)
    P = p2.shell_map_params(v, c)
    tree = cKDTree(np.array([v[t].mean(0) for t in c]))
    for row in g:
        X = row[:12].reshape(4, 3)
        d, ci = tree.query(X.mean(0))
        assert d < 1e-12
        for k in range(5):
            x = rc.BARY[k] @ X
            assert np.abs(p2.shell_eval_f(P[ci], x) - row[12 + 12 * k:15 + 12 * k]).max() < 1e-14
            assert np.abs(p2.shell_eval_df(P[ci], x) - row[15 + 12 * k:24 + 12 * k].reshape(3, 3)).max() < 1e-13


@pytest.mark.parametrize("L", [2, 3])
def test_divkgrad_blending_on_shell(L):
    g = rc.gold_p2()
    o = rc.p2_oracle("shell", L)
    x, d0 = rc.p2_inputs(o.n, 300 + L)
    M = o.matrix(form="divkgrad")
    assert rc.rel(p2.apply(o, M, x, d0, 1), g["shell_L%d_divkgrad_inner" % L]) < TOL
    if L == 2:
        assert rc.rel(p2.apply(o, M, x, d0, 15, add=True), g["shell_L2_divkgrad_add_all"]) < TOL


@pytest.mark.parametrize("i,j", [(0, 0), (0, 1), (1, 2), (2, 0), (2, 2)])
def test_epsilon_blending_on_shell(i, j):
    g = rc.gold_p2()
    o = rc.p2_oracle("shell", 2)
    x, d0 = rc.p2_inputs(o.n, 302)
    M = o.matrix(form="epsilon", comp=(i, j))
    assert rc.rel(p2.apply(o, M, x, d0, 1), g["shell_L2_epsilon_%d%d_inner" % (i, j)]) < TOL


@pytest.mark.parametrize("L", [2, 3])
def test_divkgrad_affine_on_cube(L):
    g = rc.gold_p2()
    o = rc.p2_oracle("cube", L)
    x, d0 = rc.p2_inputs(o.n, 300 + L)
    assert rc.rel(p2.apply(o, o.matrix(form="divkgrad"), x, d0, 1), g["cube_6el_L%d_divkgrad_inner" % L]) < TOL


@pytest.mark.parametrize("kind,Lc", [("cube", 2), ("shell", 1)])
def test_quadratic_transfer(kind, Lc):
    """P2toP2QuadraticProlongation = nodal interpolation, P2toP2QuadraticRestriction = its transpose (flag Inner | NeumannBoundary)"""
    g = rc.gold_p2()
    v, c = meshes.cube_6el() if kind == "cube" else meshes.spherical_shell(*rc.SHELL)
    oc, of = p2.P2Oracle(v, c, Lc, shell=False), p2.P2Oracle(v, c, Lc + 1, shell=False)
    P = p2.quadratic_prolongation(oc, of)
    xc, _ = rc.p2_inputs(oc.n, 400 + Lc)
    xf, _ = rc.p2_inputs(of.n, 500 + Lc)
    m = of.mask(5)
    assert rc.rel((P @ xc)[m], g["%s_Lc%d_prolongate" % (kind, Lc)][m]) < TOL
    m = oc.mask(5)
    assert rc.rel((P.T @ xf)[m], g["%s_Lc%d_restrict" % (kind, Lc)][m]) < TOL


@pytest.mark.parametrize("comp", [0, 1, 2])
def test_stokes_div_blocks(comp):
    """B = P2ToP1ElementwiseBlendingDiv{x,y,z}Operator and B^T = P1ToP2ElementwiseBlendingDivT{x,y,z}Operator on the shell"""
    g = rc.gold_p2()
    o = rc.p2_oracle("shell", 2)
    x, d0 = rc.p2_inputs(o.n, 302)
    B = o.div_matrix(comp)
    m = (o.doftype[: o.nv] & 1) != 0
    assert rc.rel((B @ x)[m], g["shell_L2_div_%d_inner" % comp][m]) < TOL
    m = o.mask(1)
    assert rc.rel((B.T @ d0[: o.nv])[m], g["shell_L2_divt_%d_inner" % comp][m]) < TOL
