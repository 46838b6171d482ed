"""Parity cases shared by the GPU tests (product CUDA library) and the CPU host-logic tests
(test-only emulation build).  Each case compares the C-ABI path with the oracle on seeded inputs."""
import numpy as np

from hyteg_b200 import api, meshes
from hyteg_b200.capi import All, DirichletBoundary, Inner, NeumannBoundary, FreeslipBoundary, Replace, Add
from oracle import p1_oracle as po

SMOOTH = Inner | NeumannBoundary | FreeslipBoundary
TOL = 1e-12  # north_star: apply / Jacobi / Chebyshev iterates within 1e-12 relative FP64


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


class Setup:
    def __init__(self, lib, mesh, min_level, max_level, rank=0, n_ranks=1, cell_rank=None):
        v, c = mesh
        self.st = api.PrimitiveStorage(v, c, rank, n_ranks, lib=lib, cell_rank=cell_rank if n_ranks > 1 else None)
        self.st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        self.o = po.P1Oracle(v, c, min_level, max_level)
        self.A = api.P1ConstantLaplaceOperator(self.st, min_level, max_level)
        self.min_level, self.max_level = min_level, max_level
        self.cells = self.st.local_cell_global_ids()

    def fn(self, name="f"):
        return api.P1Function(name, self.st, self.min_level, self.max_level)

    def put(self, f, x, L):
        f.upload(self.o.to_cells(x, L, self.cells), L)

    def get(self, f, L, tol=0.0):
        return self.o.from_cells(f.download(L), L, self.cells, tol=tol)

    def rand(self, L, seed):
        return np.random.default_rng(seed).standard_normal(self.o.levels[L].n)


def case_apply(lib, mesh, L, flag=Inner, update=Replace):
    s = Setup(lib, mesh, L, L)
    x, d0 = s.rand(L, 1), s.rand(L, 2)
    src, dst = s.fn("src"), s.fn("dst")
    s.put(src, x, L)
    s.put(dst, d0, L)
    s.A.apply(src, dst, L, flag, update)
    got = s.get(dst, L)
    want = po.apply(s.o, x, d0, L, flag, add=(update == Add))
    assert rel(got, want) < TOL, rel(got, want)


def case_jacobi(lib, mesh, L, omega=0.66):
    s = Setup(lib, mesh, L, L)
    x, b = s.rand(L, 3), s.rand(L, 4)
    fx, fb, ft = s.fn("x"), s.fn("b"), s.fn("tmp")
    s.put(fx, x, L)
    s.put(fb, b, L)
    sm = api.WeightedJacobiSmoother(s.st, L, L, omega)
    cur = x
    for _ in range(3):
        sm.solve(s.A, fx, fb, L)
        cur = po.smooth_jac(s.o, b, cur, omega, L, SMOOTH)
    assert rel(s.get(fx, L), cur) < TOL


def case_chebyshev(lib, mesh, L, order=3):
    s = Setup(lib, mesh, L, L)
    x, b = s.rand(L, 5), s.rand(L, 6)
    b[~s.o.mask(L, SMOOTH)] = 0
    fx, fb, ft = s.fn("x"), s.fn("b"), s.fn("t")
    # spectral radius by power iteration (start vector: seeded random)
    x0 = s.rand(L, 7)
    s.put(fx, x0, L)
    rho = api.estimateRadius(s.A, L, 20, fx, ft)
    rho_o = po.estimate_radius(s.o, x0, 20, L)
    assert abs(rho - rho_o) < 1e-10 * abs(rho_o), (rho, rho_o)
    cheb = api.ChebyshevSmoother(s.st, L, L)
    cheb.setupCoefficients(order, rho)
    co = po.cheb_coefficients(order, 0.3 * rho_o, 1.2 * rho_o)
    assert rel(cheb.coefficients(), np.array(co)) < 1e-12
    s.put(fx, x, L)
    s.put(fb, b, L)
    cur = x
    for _ in range(2):
        cheb.solve(s.A, fx, fb, L)
        cur = po.chebyshev(s.o, cur, b, co, L)
    assert rel(s.get(fx, L), cur) < TOL


def case_blas(lib, mesh, L):
    s = Setup(lib, mesh, L, L)
    a, b = s.rand(L, 8), s.rand(L, 9)
    fa, fb, fc = s.fn("a"), s.fn("b"), s.fn("c")
    s.put(fa, a, L)
    s.put(fb, b, L)
    for flag in (All, Inner, DirichletBoundary):
        m = s.o.mask(L, flag)
        assert abs(fa.dotGlobal(fb, L, flag) - a[m] @ b[m]) < 1e-12 * max(1.0, abs(a[m] @ b[m]))
        assert abs(fa.sumGlobal(L, flag) - a[m].sum()) < 1e-11
        assert fa.getMaxDoFMagnitude(L, flag) == (np.abs(a[m]).max() if m.any() else 0.0)
    fc.interpolate(0.5, L, All)
    fc.assign([2.0, -3.0], [fa, fb], L, Inner)
    want = np.full(len(a), 0.5)
    m = s.o.mask(L, Inner)
    want[m] = 2 * a[m] - 3 * b[m]
    assert rel(s.get(fc, L), want) < 1e-15
    fc.add([0.25], [fa], L, All)
    want = want + 0.25 * a
    fc.add(1.5, L, DirichletBoundary)
    want[s.o.mask(L, DirichletBoundary)] += 1.5
    assert rel(s.get(fc, L), want) < 1e-15
    fc.multElementwise([fa, fb], L, All)
    assert rel(s.get(fc, L), a * b) < 1e-15
    # interpolate of an expression reproduces the coordinates
    fc.interpolate(lambda x, y, z: 1 + 2 * x - 3 * y + 0.5 * z, L, All)
    xyz = s.o.levels[L].xyz
    assert rel(s.get(fc, L, tol=1e-14), 1 + 2 * xyz[:, 0] - 3 * xyz[:, 1] + 0.5 * xyz[:, 2]) < 1e-14


def case_transfer(lib, mesh, Lc):
    s = Setup(lib, mesh, Lc, Lc + 1)
    Lf = Lc + 1
    ff, cc = s.rand(Lf, 10), s.rand(Lc, 11)
    f = s.fn("f")
    R, P = api.P1toP1LinearRestriction(), api.P1toP1LinearProlongation()
    for flag in (SMOOTH, All):
        s.put(f, ff, Lf)
        s.put(f, cc, Lc)
        R.restrict(f, Lf, flag)
        assert rel(s.get(f, Lc, tol=1e-13), po.restrict(s.o, ff, cc, Lf, flag)) < TOL
        s.put(f, cc, Lc)
        P.prolongateAndAdd(f, Lc, flag)
        assert rel(s.get(f, Lf), po.prolongate_and_add(s.o, cc, ff, Lc, flag)) < TOL
        s.put(f, ff, Lf)
        P.prolongate(f, Lc, flag)
        want = ff.copy()
        m = s.o.mask(Lf, flag)
        want[m] = (s.o.levels[Lf].P @ cc)[m]
        assert rel(s.get(f, Lf), want) < TOL


def case_linear_kernel(lib, mesh, L):
    """P1LaplaceOperator3DTest.cpp:43-131: A*1 = 0 and A*linear = 0 on inner DoFs, discrete L2 < 2.8e-13"""
    s = Setup(lib, mesh, L, L)
    u, r, one = s.fn("u"), s.fn("r"), s.fn("one")
    one.interpolate(1.0, L, All)
    npts = one.dotGlobal(one, L, Inner)
    for expr in (lambda x, y, z: 0 * x, lambda x, y, z: 1 + 0 * x, lambda x, y, z: 42 * x, lambda x, y, z: 2 * x + 3 * y - 7 * z + 1):
        u.interpolate(expr, L, All)
        r.interpolate(0.0, L, All)
        s.A.apply(u, r, L, Inner)
        err = np.sqrt(r.dotGlobal(r, L, Inner) / max(npts, 1))
        assert err < 2.8e-13, err


def gmg_rates(lib, mesh, min_level, max_level, smoother="gs", cycles=6, pre=3, post=3, coarse_iters=200):
    """residual reduction per V-cycle of the product path and of the oracle (HyTeG ordering)"""
    s = Setup(lib, mesh, min_level, max_level)
    L = max_level
    x, b = s.fn("x"), s.fn("b")
    r = s.fn("r")
    exact = lambda X, Y, Z: np.sin(X) * np.sinh(Y) + Z
    x.interpolate(exact, L, DirichletBoundary)
    if smoother == "gs":
        sm = api.GaussSeidelSmoother(s.st)
    elif smoother == "jacobi":
        sm = api.WeightedJacobiSmoother(s.st, min_level, max_level, 0.66)
    else:
        raise ValueError(smoother)
    coarse = api.CGSolver(s.st, min_level, min_level, coarse_iters, 1e-16, 1e-16)
    gmg = api.GeometricMultigridSolver(s.st, sm, coarse, min_level, max_level, pre, post)
    res = []
    s.A.residual(x, b, r, L, Inner)
    res.append(np.sqrt(r.dotGlobal(r, L, Inner)))
    for _ in range(cycles):
        gmg.solve(s.A, x, b, L)
        s.A.residual(x, b, r, L, Inner)
        res.append(np.sqrt(r.dotGlobal(r, L, Inner)))
    # oracle
    o = s.o
    xs = {l: np.zeros(o.levels[l].n) for l in range(min_level, max_level + 1)}
    bs = {l: np.zeros(o.levels[l].n) for l in range(min_level, max_level + 1)}
    xyz = o.levels[L].xyz
    md = o.mask(L, DirichletBoundary)
    xs[L][md] = exact(xyz[md, 0], xyz[md, 1], xyz[md, 2])
    if smoother == "gs":
        osm = lambda o_, x_, b_, l_: po.smooth_sor(o_, x_, b_, 1.0, l_, SMOOTH)
    else:
        osm = lambda o_, x_, b_, l_: po.smooth_jac(o_, b_, x_, 0.66, l_, SMOOTH)
    ocs = lambda o_, x_, b_, l_: po.cg(o_, x_, b_, l_, coarse_iters)[0]
    mi = o.mask(L, Inner)
    ores = [np.linalg.norm((bs[L] - o.levels[L].A @ xs[L])[mi])]
    for _ in range(cycles):
        po.vcycle(o, xs, bs, L, min_level, osm, ocs, pre, post)
        ores.append(np.linalg.norm((bs[L] - o.levels[L].A @ xs[L])[mi]))
    return np.array(res), np.array(ores), s, x, xs[L]
