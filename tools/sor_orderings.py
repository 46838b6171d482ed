"""Experiment (CPU, oracle): V(3,3) convergence factor of candidate SOR orderings for the macro-cell interior
against the reference's lexicographic hybrid.  Interface classes (vertex/edge/face) keep the oracle's ordering."""
import sys, os, time
import numpy as np, scipy.sparse as sp
from scipy.sparse.linalg import spsolve_triangular
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import p1_oracle as po
from hyteg_b200 import meshes


def interior_sets(o, L):
    lv = o.levels[L]
    pts = lv.pts  # (n,3) x,y,z per cell-array index
    N = lv.N
    inter = (pts[:, 0] > 0) & (pts[:, 1] > 0) & (pts[:, 2] > 0) & (pts.sum(1) < N)
    return pts, inter


def make_cell_sweep(o, L, scheme, omega=1.0):
    """returns f(x, rhs) doing the cell-class update in 'scheme'"""
    lv = o.levels[L]
    A = lv.A.tocsr()
    pts, inter = interior_sets(o, L)
    N = lv.N
    idx = np.nonzero(inter)[0]
    p = pts[idx]
    x_, y_, z_ = p[:, 0], p[:, 1], p[:, 2]
    if len(idx) == 0:
        return lambda x, rhs, m: x
    if scheme == "lex":
        key = None
        colours = None
    elif scheme == "c4":
        colours = (x_ + 2 * y_ + 3 * z_) % 4
    elif scheme == "c6row":
        colours = ((y_ + 2 * z_) % 3) * 2 + (x_ % 2)
    elif scheme == "c6rowb":
        colours = ((2 * y_ + z_) % 3) * 2 + (x_ % 2)
    elif scheme == "c8":
        colours = (x_ % 2) + 2 * (y_ % 2) + 4 * (z_ % 2)
    elif scheme.startswith("ylex3") or scheme.startswith("tile3") or scheme.startswith("tile4") or scheme in ("c3liney", "c3rowx"):
        # sequential in y, 3 colours (x+2z)%3 inside a y-plane; tiles in (x,z) are block-Jacobi (tile TXxTZ)
        colours = None
    # global ids per cell
    passes = []  # list of index arrays (global ids) processed as one Jacobi-like colour step, sequentially
    tri = None
    if colours is not None:
        for c in range(colours.max() + 1):
            g = np.concatenate([lv.cell_map[k][idx[colours == c]] for k in range(len(o.C))])
            passes.append(g)
        Ad = A.diagonal()
        def f(x, rhs, m):
            x = x.copy()
            for g in passes:
                g = g[m[g]]
                r = rhs[g] - A[g] @ x
                x[g] += omega * r / Ad[g]
            return x
        return f
    # ordered (triangular) schemes: build ordering list of global ids and a "block id" for Jacobi coupling
    if scheme == "lex":
        order = np.lexsort((x_, y_, z_))
        block = np.zeros(len(idx), dtype=np.int64)
    elif scheme == "c3liney":
        order = np.lexsort((y_, x_, z_, (x_ + 2 * z_) % 3))
        block = np.zeros(len(idx), dtype=np.int64)
    elif scheme.startswith("tile3"):
        # tiles in (x,z) coloured (i + 2j) % 3; inside a tile: lines along y coloured (x+2z)%3, sequential along y
        _, TX, TZ = scheme.split("_")
        TX, TZ = int(TX), int(TZ)
        ti, tj = (x_ - 1) // TX, (z_ - 1) // TZ
        order = np.lexsort((y_, x_, z_, (x_ + 2 * z_) % 3, tj, ti, (ti + 2 * tj) % 3))
        block = np.zeros(len(idx), dtype=np.int64)
    elif scheme.startswith("tile4"):
        # tiles in (x,y,z) coloured (i + 2k + 3j) % 4; inside a tile: lines along y coloured (x+2z)%3, sequential along y
        _, TX, TY, TZ = scheme.split("_")
        TX, TY, TZ = int(TX), int(TY), int(TZ)
        ti, tk, tj = (x_ - 1) // TX, (y_ - 1) // TY, (z_ - 1) // TZ
        order = np.lexsort((y_, x_, z_, (x_ + 2 * z_) % 3, tk, tj, ti, (ti + 2 * tk + 3 * tj) % 4))
        block = np.zeros(len(idx), dtype=np.int64)
    elif scheme == "c3rowx":
        order = np.lexsort((x_, y_, z_, (y_ + 2 * z_) % 3))
        block = np.zeros(len(idx), dtype=np.int64)
    else:
        _, TX, TZ = scheme.split("_")
        TX, TZ = int(TX), int(TZ)
        block = ((x_ - 1) // TX) * 1000 + ((z_ - 1) // TZ)
        col = (x_ + 2 * z_) % 3
        order = np.lexsort((col, y_, block))  # block-major, then y, then colour
    G, B = [], []
    for k in range(len(o.C)):
        G.append(lv.cell_map[k][idx[order]])
        B.append(block[order] + 1000000 * k)
    G = np.concatenate(G); B = np.concatenate(B)
    n = len(G)
    Ass = A[G][:, G].tocoo()
    same = B[Ass.row] == B[Ass.col]
    lowmask = same & (Ass.row > Ass.col)
    Lm = sp.csr_matrix((Ass.data[lowmask], (Ass.row[lowmask], Ass.col[lowmask])), shape=(n, n))
    D = A.diagonal()[G]
    M = (sp.diags(D / omega) + Lm).tocsr()
    def f(x, rhs, m):
        assert m[G].all()
        x = x.copy()
        # r = rhs - A x (old) + (D/omega + L) x_old  => solve M x_new = rhs - (A - M) x_old
        r = rhs[G] - A[G] @ x + M @ x[G]
        x[G] = spsolve_triangular(M, r, lower=True)
        return x
    return f


def smoother_factory(o, scheme):
    cache = {}
    def sm(o_, x, b, L):
        lv = o.levels[L]
        m = o.mask(L, po.SMOOTH_FLAG)
        # interface classes in the oracle's (reference) ordering
        x = sor_classes(o, x, b, 1.0, L, m, (0, 1, 2))
        if L not in cache:
            cache[L] = make_cell_sweep(o, L, scheme)
        return cache[L](x, b, m)
    return sm


def sor_classes(o, x, rhs, omega, L, m, classes):
    lv = o.levels[L]
    A = lv.A
    indptr, indices, data = A.indptr, A.indices, A.data
    x = x.copy()
    order = po._lex_order(o, L)
    prim = lv.prim_of
    for d in classes:
        snap = x.copy()
        for g in order[d]:
            if not m[g]:
                continue
            s = rhs[g]; dg = 0.0
            for k in range(indptr[g], indptr[g + 1]):
                j = indices[k]
                if j == g: dg = data[k]
                elif lv.dim[j] == d and prim[j] != prim[g]: s -= data[k] * snap[j]
                else: s -= data[k] * x[j]
            x[g] = (1.0 - omega) * x[g] + omega * s / dg
    return x


def run(mesh, Lmax, scheme, cycles=int(os.environ.get("CYCLES", 8))):
    v, c = mesh
    o = po.P1Oracle(v, c, 0, Lmax)
    rng = np.random.default_rng(1)
    lv = o.levels[Lmax]
    m = o.mask(Lmax, po.INNER)
    x = {L: np.zeros(o.levels[L].n) for L in range(Lmax + 1)}
    b = {L: np.zeros(o.levels[L].n) for L in range(Lmax + 1)}
    x[Lmax][m] = rng.random(m.sum())
    sm = smoother_factory(o, scheme)
    coarse = lambda o_, x_, b_, L: po.cg(o, x_, b_, L, 1000)[0]
    res = []
    for k in range(cycles + 1):
        r = (b[Lmax] - lv.A @ x[Lmax])[m]
        res.append(np.sqrt(r @ r))
        if k < cycles:
            po.vcycle(o, x, b, Lmax, 0, sm, coarse)
    res = np.array(res)
    return res[1:] / res[:-1]


if __name__ == "__main__":
    Lmax = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    mesh = meshes.cube_6el() if len(sys.argv) < 4 else meshes.golden(sys.argv[3])
    for scheme in sys.argv[2].split(","):
        t = time.time()
        f = run(mesh, Lmax, scheme)
        print(scheme, "factors", np.array2string(f, precision=4), "asym", (f[-1] * f[-2]) ** 0.5, "%.1fs" % (time.time() - t), flush=True)
