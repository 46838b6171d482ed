"""P2 parity cases shared by the GPU tests (product CUDA library) and the CPU host-logic tests (test-only emulation build): the C-ABI
path against tests/golden/ref_p2.npz (outputs of the reference itself) and against oracle/p2_oracle.py."""
import numpy as np

from hyteg_b200 import api, meshes
from hyteg_b200.capi import Add, All, Inner, Replace
from oracle import p2_oracle as p2
import ref_cases as rc

TOL = 1e-12


class P2Setup:
    def __init__(self, lib, kind, L):
        self.o = rc.p2_oracle(kind, L)
        v, c = (meshes.spherical_shell(*rc.SHELL) if kind == "shell" else meshes.cube_6el())
        self.st = api.PrimitiveStorage(v, c, lib=lib)
        self.st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        if kind == "shell":
            api.setIcosahedralShellMap(self.st)
        self.L = L

    def fn(self, name, x=None):
        f = api.P2Function(name, self.st, self.L, self.L)
        if x is not None:
            v, e = self.o.to_cells(x)
            f.upload(v, e, self.L)
        return f

    def get(self, f, tol=0.0):
        v, e = f.download(self.L)
        return self.o.from_cells(v, e, tol=tol)


def k_xyz(x, y, z):
    return p2.coefficient_k(np.stack([x, y, z], axis=-1))


def case_shell_map(lib):
    """device evalF / evalDF and the host classification against the reference's IcosahedralShellMap samples"""
    from scipy.spatial import cKDTree
    for nt, nr in ((3, 2), (2, 3)):
        g = rc.gold_p2()["shellmap_%d_%d" % (nt, nr)]
        v, c = meshes.spherical_shell(nt, nr, 0.5, 1.0)
        st = api.PrimitiveStorage(v, c, lib=lib)
        api.setIcosahedralShellMap(st)
        tree = cKDTree(np.array([v[t].mean(0) for t in c]))
        for row in g[:: max(1, len(g) // 60)]:
            X = row[:12].reshape(4, 3)
            d, ci = tree.query(X.mean(0))
            F, DF = api.evalMap(st, 1, int(ci), rc.BARY @ X)
            for k in range(5):
                assert np.abs(F[k] - row[12 + 12 * k:15 + 12 * k]).max() < 1e-14
                assert np.abs(DF[k] - row[15 + 12 * k:24 + 12 * k].reshape(3, 3)).max() < 1e-13


def case_divkgrad(lib, kind, L, mode="sampled"):
    """P2ElementwiseBlendingDivKGradOperator::apply against the reference's output (k through the callback path) and, with k as a
    P2 function (the hyteg_operators interface), against the oracle"""
    g = rc.gold_p2()
    s = P2Setup(lib, kind, L)
    x, d0 = rc.p2_inputs(s.o.n, 300 + L)
    src, dst = s.fn("src", x), s.fn("dst", d0)
    if mode == "sampled":
        A = api.P2ElementwiseBlendingDivKGradOperator(s.st, L, L, k_xyz)
        A.apply(src, dst, L, Inner, Replace)
        name = ("shell_L%d_divkgrad_inner" if kind == "shell" else "cube_6el_L%d_divkgrad_inner") % L
        assert rc.rel(s.get(dst, tol=1e-13), g[name]) < TOL
        if kind == "shell" and L == 2:
            dst = s.fn("dst", d0)
            A.apply(src, dst, L, All, Add)
            assert rc.rel(s.get(dst, tol=1e-13), g["shell_L2_divkgrad_add_all"]) < TOL
    else:
        kvec = p2.coefficient_k(s.o.dof_physical_coordinates())
        kf = s.fn("k", kvec)
        A = api.P2ElementwiseBlendingDivKGradOperator(s.st, L, L, kf)
        A.apply(src, dst, L, Inner, Replace)
        want = p2.apply(s.o, s.o.matrix(form="divkgrad", kvec=kvec), x, d0, Inner)
        assert rc.rel(s.get(dst, tol=1e-13), want) < TOL


def case_epsilon(lib, i, j):
    g = rc.gold_p2()
    s = P2Setup(lib, "shell", 2)
    x, d0 = rc.p2_inputs(s.o.n, 302)
    src, dst = s.fn("src", x), s.fn("dst", d0)
    A = api.P2ElementwiseBlendingEpsilonBlock(s.st, 2, 2, i, j, k_xyz)
    A.apply(src, dst, 2, Inner, Replace)
    assert rc.rel(s.get(dst, tol=1e-13), g["shell_L2_epsilon_%d%d_inner" % (i, j)]) < TOL


def case_blas(lib):
    s = P2Setup(lib, "shell", 2)
    a, b = rc.p2_inputs(s.o.n, 7)
    fa, fb, fc = s.fn("a", a), s.fn("b", b), s.fn("c")
    for flag in (All, Inner):
        m = s.o.mask(flag)
        assert abs(fa.dotGlobal(fb, 2, flag) - a[m] @ b[m]) < 1e-11 * abs(a[m] @ b[m])
    fc.interpolate(0.5, 2, All)
    fc.assign([2.0, -3.0], [fa, fb], 2, Inner)
    want = np.full(s.o.n, 0.5)
    m = s.o.mask(Inner)
    want[m] = 2 * a[m] - 3 * b[m]
    assert rc.rel(s.get(fc), want) < 1e-15
    fc.add([0.25], [fa], 2, All)
    assert rc.rel(s.get(fc), want + 0.25 * a) < 1e-15


def case_transfer(lib, kind, Lc):
    """P2toP2QuadraticRestriction / Prolongation against the reference's output (flag Inner | NeumannBoundary) and prolongateAndAdd
    against the oracle"""
    g = rc.gold_p2()
    v, c = meshes.cube_6el() if kind == "cube" else meshes.spherical_shell(*rc.SHELL)
    oc, of = p2.P2Oracle(v, c, Lc, shell=False), p2.P2Oracle(v, c, Lc + 1, shell=False)
    st = api.PrimitiveStorage(v, c, lib=lib)
    st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    xc, _ = rc.p2_inputs(oc.n, 400 + Lc)
    xf, _ = rc.p2_inputs(of.n, 500 + Lc)
    f = api.P2Function("f", st, Lc, Lc + 1)
    R, P = api.P2toP2QuadraticRestriction(), api.P2toP2QuadraticProlongation()

    def put(o, x, L):
        vv, ee = o.to_cells(x)
        f.v.upload(vv, L, All)
        f.lib.check(f.lib.dll.hb_p2function_upload_edges(f.h, L, api.as_dp(api.f64(ee))))

    def get(o, L):
        ee = np.zeros(f.edge_size(L))
        f.lib.check(f.lib.dll.hb_p2function_download_edges(f.h, L, api.as_dp(ee)))
        return o.from_cells(f.v.download(L), ee, tol=1e-13)

    put(oc, xc, Lc)
    put(of, np.zeros(of.n), Lc + 1)
    P.prolongate(f, Lc, 5)
    m = of.mask(5)
    got = get(of, Lc + 1)
    assert rc.rel(got[m], g["%s_Lc%d_prolongate" % (kind, Lc)][m]) < TOL
    assert np.abs(got[~m]).max() == 0.0
    put(of, xf, Lc + 1)
    P.prolongateAndAdd(f, Lc, 5)
    want = xf.copy()
    want[m] += (p2.quadratic_prolongation(oc, of) @ xc)[m]
    assert rc.rel(get(of, Lc + 1), want) < TOL
    put(of, xf, Lc + 1)
    put(oc, xc, Lc)
    R.restrict(f, Lc + 1, 5)
    m = oc.mask(5)
    got = get(oc, Lc)
    assert rc.rel(got[m], g["%s_Lc%d_restrict" % (kind, Lc)][m]) < TOL
    assert np.array_equal(got[~m], xc[~m])  # DoFs of unselected primitives keep their values


def case_div_blocks(lib, comp):
    g = rc.gold_p2()
    s = P2Setup(lib, "shell", 2)
    x, d0 = rc.p2_inputs(s.o.n, 302)
    u = s.fn("u", x)
    q = api.P1Function("q", s.st, 2, 2)
    api.P2ToP1ElementwiseBlendingDivOperator(s.st, 2, 2, comp).apply(u, q, 2, Inner)
    got = s.o.p1.from_cells(q.download(2), 2, tol=1e-13)
    m = (s.o.doftype[: s.o.nv] & 1) != 0
    assert rc.rel(got[m], g["shell_L2_div_%d_inner" % comp][m]) < TOL
    assert np.abs(got[~m]).max() == 0.0
    q.upload(s.o.p1.to_cells(d0[: s.o.nv], 2), 2)
    w = s.fn("w")
    api.P1ToP2ElementwiseBlendingDivTOperator(s.st, 2, 2, comp).apply(q, w, 2, Inner)
    m = s.o.mask(Inner)
    assert rc.rel(s.get(w, tol=1e-13)[m], g["shell_L2_divt_%d_inner" % comp][m]) < TOL


def case_stokes(lib, fused=True):
    """the fused vector epsilon kernel and the Taylor-Hood composite against the blocks of the oracle (each block pinned to the
    reference: tests/test_p2_oracle_vs_reference.py)"""
    s = P2Setup(lib, "shell", 2)
    o = s.o
    rng = np.random.default_rng(11)
    U = [rng.random(o.n) - 0.5 for _ in range(3)]
    pv = rng.random(o.nv) - 0.5
    src, dst = api.P2P1TaylorHoodFunction("s", s.st, 2, 2), api.P2P1TaylorHoodFunction("d", s.st, 2, 2)
    for c in range(3):
        v, e = o.to_cells(U[c])
        src.uvw[c].upload(v, e, 2)
    src.p.upload(o.p1.to_cells(pv, 2), 2)
    S = api.P2P1ElementwiseBlendingEpsilonStokesOperator(s.st, 2, 2, k_xyz, fused=fused)
    S.apply(src, dst, 2, Inner)
    m = o.mask(Inner)
    B = [o.div_matrix(c) for c in range(3)]
    for c in range(3):
        want = sum(o.matrix(form="epsilon", comp=(c, d)) @ U[d] for d in range(3)) + B[c].T @ pv
        got = s.get(dst.uvw[c], tol=1e-12)
        assert rc.rel(got[m], want[m]) < 1e-11
    wantp = sum(B[c] @ U[c] for c in range(3))
    gotp = o.p1.from_cells(dst.p.download(2), 2, tol=1e-12)
    mp = (o.doftype[: o.nv] & 1) != 0
    assert rc.rel(gotp[mp], wantp[mp]) < 1e-11


def case_chebyshev_gmg(lib, cycles=5):
    """Chebyshev smoother (order 3, inverse diagonal) and the Chebyshev-smoothed V(2,2) GMG, levels 1..3 on the blended shell, against
    the REFERENCE's own run of GeometricMultigridSolver< P2ElementwiseOperator< p2_div_k_grad_blending_q4 > > (tests/golden/ref_p2.npz)"""
    g = rc.gold_p2()
    v, c = meshes.spherical_shell(*rc.SHELL)
    o = rc.p2_oracle("shell", 3)
    st = api.PrimitiveStorage(v, c, lib=lib)
    st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    api.setIcosahedralShellMap(st)
    rng = np.random.default_rng(9)
    x0, b, ev = rng.random(o.n), rng.random(o.n) - 0.5, rng.random(o.n)

    def fn(name, val=None):
        f = api.P2Function(name, st, 1, 3)
        if val is not None:
            vv, ee = o.to_cells(val)
            f.upload(vv, ee, 3)
        return f

    def get(f):
        vv, ee = f.download(3)
        return o.from_cells(vv, ee, tol=1e-12)

    A = api.P2ElementwiseBlendingDivKGradOperator(st, 1, 3, k_xyz)
    x, rhs, fev = fn("x", x0), fn("b", b), fn("ev", ev)
    S = api.P2ChebyshevGMG(st, A, 1, 3, 3, fev, power_iterations=20, pre=2, post=2)
    assert abs(S.spectral_radius - g["shell_L3_rho20"][0]) < 1e-10 * S.spectral_radius
    S.smooth(x, rhs, 3)
    S.smooth(x, rhs, 3)
    assert rc.rel(get(x), g["shell_L3_cheb3_2sweeps"]) < 1e-11
    x = fn("x", x0)
    r = fn("r")
    m = o.mask(Inner)
    res = []
    for k in range(cycles + 1):
        A.apply(x, r, 3, Inner)
        r.assign([1.0, -1.0], [rhs, r], 3, Inner)
        res.append(r.dotGlobal(r, 3, Inner) / m.sum())
        if k < cycles:
            S.solve(x, rhs, 3)
    want = g["shell_gmg_cheb_res2"]
    assert np.allclose(res, want[: cycles + 1], rtol=1e-7), (res, want)
    if cycles == 5:
        assert rc.rel(get(x), g["shell_gmg_cheb_x"]) < 1e-9
