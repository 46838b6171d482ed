"""Pins the oracle (oracle/p1_oracle.py) against the reference: its known-answer tests and, when oracle/_ref was
built from /root/reference, the reference's own generated macro-cell kernels."""
import ctypes as C
import os

import numpy as np
import pytest

from hyteg_b200 import meshes
from oracle import p1_oracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libhyteg_ref_kernels.so")
MESHES = ["cube_6el", "regular_octahedron_8el", "pyramid_tilted_4el", "three_tets_with_two_joint_faces", "tet_tilted_1el", "cube_24el", "pyramid_4el"]
DIRS = np.array([[0, 0, 0], [-1, 0, 0], [1, 0, 0], [0, 1, 0], [0, -1, 0], [-1, 1, 0], [1, -1, 0], [0, 0, 1], [-1, 0, 1], [0, -1, 1],
                 [1, -1, 1], [0, 0, -1], [0, 1, -1], [1, 0, -1], [-1, 1, -1]])


@pytest.mark.parametrize("name", MESHES)
@pytest.mark.parametrize("L", [2, 3])
def test_constants_and_linears_in_kernel(name, L):
    """P1LaplaceOperator3DTest.cpp:43-131: A*1 = 0, A*linear = 0 on inner DoFs, discrete L2 < 2.8e-13"""
    v, c = meshes.golden(name)
    o = po.P1Oracle(v, c, L, L)
    lv = o.levels[L]
    m = o.mask(L, po.INNER)
    for u in (np.ones(lv.n), 42 * lv.xyz[:, 0], 2 * lv.xyz[:, 0] + 3 * lv.xyz[:, 1] - 7 * lv.xyz[:, 2] + 1):
        r = po.apply(o, u, np.zeros(lv.n), L, po.INNER)
        assert np.sqrt((r[m] ** 2).sum() / max(m.sum(), 1)) < 2.8e-13


def cell_stencil(o, L, cell):
    """interior 15-point stencil of a macro-cell read off the assembled matrix (needs L >= 2)"""
    lv = o.levels[L]
    N = lv.N
    p = np.array([1, 1, 1]) if N >= 4 else None
    assert p is not None
    row = lv.cell_map[cell][po.cell_index(N, *p)]
    w = np.zeros(15)
    for d in range(15):
        q = p + DIRS[d]
        w[d] = lv.A[row, lv.cell_map[cell][po.cell_index(N, *q)]]
    return w


@pytest.mark.parametrize("name", ["cube_6el", "tet_tilted_1el", "regular_octahedron_8el"])
def test_stencil_symmetry_rowsum_scaling(name):
    """VertexDoFStencilAssemblyTest.cpp:64-238: row sum 0, W=E, N=S, NW=SE, TC=BC, TW=BE, TS=BN, TSE=BNW, x2 per level"""
    v, c = meshes.golden(name)
    o = po.P1Oracle(v, c, 2, 3)
    for cell in range(len(c)):
        w2, w3 = cell_stencil(o, 2, cell), cell_stencil(o, 3, cell)
        for w in (w2, w3):
            assert abs(w.sum()) < 1e-13
            for a, b in ((1, 2), (3, 4), (5, 6), (7, 11), (8, 13), (9, 12), (10, 14)):
                assert abs(w[a] - w[b]) < 1e-13
        # h -> h/2 scales the 3D stiffness stencil by 1/2
        assert np.allclose(w3, 0.5 * w2, atol=1e-13)


def test_gmg_convergence_bound():
    """P1GMG3DConvergenceTest.cpp:148: V(3,3) + GS, regular_octahedron_8el, levels 0..3: squared residual ratio < 3.2e-2"""
    v, c = meshes.golden("regular_octahedron_8el")
    o = po.P1Oracle(v, c, 0, 3)
    L = 3
    xs = {l: np.zeros(o.levels[l].n) for l in range(4)}
    bs = {l: np.zeros(o.levels[l].n) for l in range(4)}
    xyz = o.levels[L].xyz
    md, mi = o.mask(L, po.DIRICHLET), o.mask(L, po.INNER)
    xs[L][md] = np.sin(xyz[md, 0]) * np.sinh(xyz[md, 1])
    sm = lambda o_, x_, b_, l_: po.smooth_sor(o_, x_, b_, 1.0, l_, po.SMOOTH_FLAG)
    cs = lambda o_, x_, b_, l_: po.cg(o_, x_, b_, l_, 100)[0]
    res = [np.linalg.norm((bs[L] - o.levels[L].A @ xs[L])[mi])]
    for _ in range(4):
        po.vcycle(o, xs, bs, L, 0, sm, cs, 3, 3)
        res.append(np.linalg.norm((bs[L] - o.levels[L].A @ xs[L])[mi]))
    res = np.array(res)
    assert ((res[1:] / res[:-1]) ** 2 < 3.2e-2).all(), res


# ---- against the reference's generated kernels (oracle/_ref) ------------------------------------
needs_ref = pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref not built (needs /root/reference)")


def ref():
    dll = C.CDLL(REF_SO)
    dp = C.POINTER(C.c_double)
    dll.ref_apply_macrocell.argtypes = [dp, dp, C.c_int, dp, C.c_int]
    dll.ref_sor_macrocell.argtypes = [dp, dp, C.c_int, dp, C.c_double, C.c_int]
    dll.ref_restrict_macrocell.argtypes = [dp, dp, C.c_int, dp]
    dll.ref_prolongate_macrocell.argtypes = [dp, dp, C.c_int, dp]
    return dll, dp


@needs_ref
@pytest.mark.parametrize("L", [2, 3, 4])
def test_oracle_matches_reference_apply_kernel(L):
    """single macro-tet: oracle A restricted to the interior rows == apply_3D_macrocell_vertexdof_to_vertexdof_replace"""
    dll, dp = ref()
    v, c = meshes.golden("tet_tilted_1el")
    o = po.P1Oracle(v, c, L, L)
    lv = o.levels[L]
    w = cell_stencil(o, L, 0)
    x = np.random.default_rng(0).standard_normal(lv.n)
    src = o.to_cells(x, L)
    dst = np.zeros_like(src)
    dll.ref_apply_macrocell(dst.ctypes.data_as(dp), src.ctypes.data_as(dp), L, w.ctypes.data_as(dp), 0)
    want = o.from_cells(dst, L)
    got = lv.A @ x
    m = lv.dim == 3
    assert m.sum() == (lv.N - 1) * (lv.N - 2) * (lv.N - 3) // 6
    assert np.abs(got[m] - want[m]).max() < 1e-12 * np.abs(want[m]).max()


@needs_ref
@pytest.mark.parametrize("backwards", [0, 1])
def test_oracle_matches_reference_sor_kernel(backwards):
    """single macro-tet with Dirichlet boundary: the oracle's smooth_sor (cell class) == sor_3D_macrocell_P1[_backwards]"""
    dll, dp = ref()
    L = 3
    v, c = meshes.golden("tet_tilted_1el")
    o = po.P1Oracle(v, c, L, L)
    lv = o.levels[L]
    w = cell_stencil(o, L, 0)
    rng = np.random.default_rng(1)
    x, b = rng.standard_normal(lv.n), rng.standard_normal(lv.n)
    dst, rhs = o.to_cells(x, L).copy(), o.to_cells(b, L)
    dll.ref_sor_macrocell(dst.ctypes.data_as(dp), rhs.ctypes.data_as(dp), L, w.ctypes.data_as(dp), 1.3, backwards)
    want = o.from_cells(dst, L)
    got = po.smooth_sor(o, x, b, 1.3, L, po.INNER, backwards=bool(backwards))
    assert np.abs(got - want).max() < 1e-12 * np.abs(want).max()


@needs_ref
@pytest.mark.parametrize("Lc", [1, 2, 3])
def test_oracle_matches_reference_transfer_kernels(Lc):
    """single macro-tet (all neighbour counts 1): P and P^T of the oracle == the generated push/pull kernels"""
    dll, dp = ref()
    v, c = meshes.golden("tet_tilted_1el")
    o = po.P1Oracle(v, c, Lc, Lc + 1)
    P = o.levels[Lc + 1].P
    rng = np.random.default_rng(2)
    xc, xf = rng.standard_normal(o.levels[Lc].n), rng.standard_normal(o.levels[Lc + 1].n)
    nn = np.ones(14)
    fine = np.zeros(o.levels[Lc + 1].n)
    coarse = o.to_cells(xc, Lc)
    dll.ref_prolongate_macrocell(coarse.ctypes.data_as(dp), fine.ctypes.data_as(dp), Lc, nn.ctypes.data_as(dp))
    assert np.abs(o.from_cells(fine, Lc + 1) - P @ xc).max() < 1e-13
    cdst = np.zeros(o.levels[Lc].n)
    fsrc = o.to_cells(xf, Lc + 1)
    dll.ref_restrict_macrocell(cdst.ctypes.data_as(dp), fsrc.ctypes.data_as(dp), Lc, nn.ctypes.data_as(dp))
    assert np.abs(o.from_cells(cdst, Lc) - P.T @ xf).max() < 1e-12
