"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy / scipy) of the reference's P2 elementwise path.

Follows, in the reference tree:
  * P2Function = VertexDoFFunction + EdgeDoFFunction (src/hyteg/p2functionspace/P2Function.hpp:76-77); edge DoFs of a macro-cell in 7
    orientation arrays X, Y, Z, XY, XZ, YZ (tetrahedral arrays of width 2^L) and XYZ (width 2^L - 1)
    (src/hyteg/edgedofspace/EdgeDoFIndexing.hpp:922-985, EdgeDoFOrientation.hpp:29-39; end points of an edge DoF :170-209);
  * the six micro-cell types and their vertices (src/hyteg/volumedofspace/CellDoFIndexing.hpp:155-198) and the local numbering of the ten
    P2 DoFs of a micro-tet: 4 vertices, then the edges (2,3) (1,3) (1,2) (0,3) (0,2) (0,1) (EdgeDoFIndexing.hpp:1365-1396);
  * P2ElementwiseOperator::gemv (src/hyteg/elementwiseoperators/P2ElementwiseOperator.cpp:109-320): sum of the element matrices, DoFs of
    primitives whose boundary type is not selected stay untouched;
  * the element matrices of forms/form_hyteg_generated/p2/p2_div_k_grad_blending_q4.cpp (Xiao-Gimbutas rule, 11 points, degree 4),
    p2_epsilonvar_blending_q2.cpp (4 points, degree 2) and the P2 Laplace form, restated from their definitions
        a_ij = sum_q w_q k(F(x_q)) |det J_aff| |det DF(x_q)| (G grad phi_i) . (G grad phi_j),  G = (DF J_aff)^-T,
    with the quadrature points and weights read off the generated files (constant tables);
  * IcosahedralShellMap::evalF / evalDF / classifyVertices (src/hyteg/geometry/IcosahedralShellMap.hpp:109-239, 529-776).

Pinned against the reference itself (oracle/_ref/ref_dump p2apply / shellmap, tests/golden/ref_p2.npz).
"""
import numpy as np
import scipy.sparse as sp

from . import p1_oracle as po

# end points (offsets from the edge DoF's logical index) per orientation X, Y, Z, XY, XZ, YZ, XYZ
EDGE_ENDS = np.array([[(0, 0, 0), (1, 0, 0)], [(0, 0, 0), (0, 1, 0)], [(0, 0, 0), (0, 0, 1)], [(1, 0, 0), (0, 1, 0)],
                      [(1, 0, 0), (0, 0, 1)], [(0, 1, 0), (0, 0, 1)], [(0, 1, 0), (1, 0, 1)]], dtype=np.int64)
# micro-cell types WHITE_UP, BLUE_UP, GREEN_UP, WHITE_DOWN, BLUE_DOWN, GREEN_DOWN: vertex offsets from the cell's logical index
CELL_TYPES = np.array([[(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1)], [(1, 0, 0), (1, 1, 0), (0, 1, 0), (1, 0, 1)],
                       [(1, 0, 0), (0, 1, 0), (1, 0, 1), (0, 0, 1)], [(1, 1, 0), (1, 1, 1), (0, 1, 1), (1, 0, 1)],
                       [(1, 0, 1), (0, 1, 1), (0, 0, 1), (0, 1, 0)], [(0, 1, 0), (1, 1, 0), (1, 0, 1), (0, 1, 1)]], dtype=np.int64)
CELL_TYPE_SLACK = (1, 2, 2, 3, 2, 2)  # a micro-cell of the type exists at (x,y,z) iff x + y + z <= N - slack
EDGE_PAIRS = ((2, 3), (1, 3), (1, 2), (0, 3), (0, 2), (0, 1))

# quadrature rules in barycentric coordinates of the micro-tet's vertices 0..3 (weights sum to 1/6)
XG4_POINTS = np.array([
    [0.17469405869723054, 0.040490506727590428, 0.01356070187980288, 0.77125473269537614],
    [0.081404918402859225, 0.75250850700965499, 0.068099370938206658, 0.097987203649279112],
    [0.30444840243449711, 0.24027666492807259, 0.12680172591539199, 0.32847320672203839],
    [0.74122888209362259, 0.067223294893383398, 0.035183929773598722, 0.15636389323939531],
    [0.053341239535745266, 0.41926631387951302, 0.047781435559086663, 0.47961101102565512],
    [0.43295349048135578, 0.45076587609127677, 0.059456616299433829, 0.056824017127933668],
    [0.53800720391618573, 0.12941137378891041, 0.33019041483746447, 0.0023910074574393651],
    [0.008991260093335729, 0.1215419913339278, 0.30649398842969028, 0.56297276014304609],
    [0.10660417256199362, 0.097204644587583267, 0.68439041545304002, 0.11180076739738309],
    [0.32923295974264682, 0.029569495206479609, 0.31790356021339461, 0.32329398483747901],
    [0.10384411641099309, 0.43271023904776862, 0.35382323920929709, 0.1096224053319412]])
XG4_WEIGHTS = np.array([0.006541848487473325, 0.0092122281926561474, 0.032235135341605742, 0.0092322998119293929, 0.0099888641910932524,
                        0.011578327656272558, 0.012693785874259723, 0.013237780011337548, 0.017744672359248346, 0.018372372071416277,
                        0.025829352669374347])
_a, _b = 0.13819660112501059, 0.58541019662496829
VR1_POINTS = np.array([[_a, _a, _a, _b], [_a, _a, _b, _a], [_a, _b, _a, _a], [_b, _a, _a, _a]])
VR1_WEIGHTS = np.full(4, 0.041666666666666657)

TOL_RAY = 1e-14  # IcosahedralShellMap.hpp:597
GRAD_LAMBDA = np.array([[-1.0, -1.0, -1.0], [1, 0, 0], [0, 1, 0], [0, 0, 1]])


def p2_shape(lam):
    """values [10] and reference gradients [10,3] of the P2 basis at barycentric point lam[4]"""
    phi = np.zeros(10)
    g = np.zeros((10, 3))
    for i in range(4):
        phi[i] = lam[i] * (2 * lam[i] - 1)
        g[i] = (4 * lam[i] - 1) * GRAD_LAMBDA[i]
    for k, (a, b) in enumerate(EDGE_PAIRS):
        phi[4 + k] = 4 * lam[a] * lam[b]
        g[4 + k] = 4 * (lam[a] * GRAD_LAMBDA[b] + lam[b] * GRAD_LAMBDA[a])
    return phi, g


def coefficient_k(p):
    """the variable coefficient of the fixtures (oracle/ref_dump_p2.cpp: coefficientK)"""
    return 2.0 + np.sin(3.0 * p[..., 0]) * np.cos(2.0 * p[..., 1]) + p[..., 2] ** 2


# ---- IcosahedralShellMap ---------------------------------------------------------------------------------------------

def _classify_tet(X):
    rad = np.linalg.norm(X, axis=1)
    eps = (rad.max() - rad.min()) * 0.001
    n_outer = int(((rad.max() - rad) < eps).sum())
    return {1: "out", 2: "skew", 3: "in"}[n_outer], rad


def shell_map_params(V, C):
    """per macro-cell [ray(3), ref(3), thr(3), for(3), radRef, radRay] (IcosahedralShellMap::classifyVertices incl. the
    non-skew neighbour search of findNonSkewTetInPrism, IcosahedralShellMap.hpp:608-776)"""
    V = np.asarray(V, dtype=np.float64)
    C = np.asarray(C)
    face_cells = {}
    for c, t in enumerate(C):
        for f in ((0, 1, 2), (0, 1, 3), (0, 2, 3), (1, 2, 3)):
            face_cells.setdefault(tuple(sorted(int(t[i]) for i in f)), []).append(c)

    def classify(c, depth=0):
        X = V[C[c]]
        kind, rad = _classify_tet(X)
        ref = ray = None
        for k in range(4):
            for j in range(k + 1, 4):
                cr = np.cross(X[k], X[j])
                if (np.abs(cr) / (rad[k] * rad[j]) > TOL_RAY).any():
                    continue
                ref, ray = k, j
                break
            if ref is not None:
                break
        assert ref is not None, "no vertex pair on a radial ray"
        if kind == "skew":
            assert depth < 4
            for f in ((0, 1, 2), (0, 1, 3), (0, 2, 3), (1, 2, 3)):
                if ref in f and ray in f:
                    continue
                for other in face_cells[tuple(sorted(int(C[c][i]) for i in f))]:
                    if other != c:
                        return classify(other, depth + 1)
            raise AssertionError("could not find a non-skew tet in the prism")
        thr = fr = None
        for k in range(4):
            if k != ref and k != ray:
                thr = k
        for k in range(4):
            if k != ref and k != ray and k != thr:
                fr = k
        if (kind == "out" and rad[ref] < rad[ray]) or (kind == "in" and rad[ref] > rad[ray]):
            ref, ray = ray, ref
        return np.concatenate([X[ray], X[ref], X[thr], X[fr], [np.linalg.norm(X[ref]), np.linalg.norm(X[ray])]])

    return np.array([classify(c) for c in range(len(C))])


def shell_eval_f(P, x):
    """IcosahedralShellMap::evalF (IcosahedralShellMap.hpp:109-147); P [...,14], x [...,3]"""
    ray, ref, thr, fr = P[..., 0:3], P[..., 3:6], P[..., 6:9], P[..., 9:12]
    rad_ref, rad_ray = P[..., 12], P[..., 13]
    det3 = lambda a, b, c: np.einsum("...i,...i->...", a, np.cross(b, c))
    vol_t = det3(ref - ray, thr - ray, fr - ray)
    vol_x = det3(x - ray, thr - ray, fr - ray)
    bary = np.abs(vol_x / vol_t)
    old = np.linalg.norm(x, axis=-1)
    new = rad_ray + bary * (rad_ref - rad_ray)
    return x * (new / old)[..., None]


def shell_eval_df(P, x):
    """Jacobian of evalF, restated from the definition F(x) = x r(x) / |x| with r affine in x:
    DF = (r / |x|) I + x (grad r)^T / |x| - r x x^T / |x|^3   (the reference's evalDF, :169-239, is the expanded form)"""
    ray, ref, thr, fr = P[..., 0:3], P[..., 3:6], P[..., 6:9], P[..., 9:12]
    rad_ref, rad_ray = P[..., 12], P[..., 13]
    det3 = lambda a, b, c: np.einsum("...i,...i->...", a, np.cross(b, c))
    n = np.cross(thr - ray, fr - ray)  # vol_x = (x - ray) . n
    vol_t = det3(ref - ray, thr - ray, fr - ray)
    vol_x = np.einsum("...i,...i->...", x - ray, n)
    # (inside the macro-cell vol_x / vol_t >= 0; like the reference's evalDF this differentiates the ratio itself, not its modulus)
    r = rad_ray + (vol_x / vol_t) * (rad_ref - rad_ray)
    grad_r = ((rad_ref - rad_ray) / vol_t)[..., None] * n
    nx = np.linalg.norm(x, axis=-1)
    eye = np.eye(3)
    return (r / nx)[..., None, None] * eye + x[..., :, None] * grad_r[..., None, :] / nx[..., None, None] \
        - (r / nx ** 3)[..., None, None] * x[..., :, None] * x[..., None, :]


# ---- P2 function space -----------------------------------------------------------------------------------------------

def tet_size(w):
    return w * (w + 1) * (w + 2) // 6


class P2Oracle:
    """Unique-DoF view of a P2 function space on one level: vertex DoFs 0..nv-1 (the P1 oracle's numbering), edge DoFs nv..nv+ne-1."""

    def __init__(self, vertices, cells, level, shell=False):
        self.L = level
        self.p1 = po.P1Oracle(vertices, cells, level, level)
        self.V, self.C = self.p1.V, self.p1.C
        lv = self.p1.levels[level]
        self.lv = lv
        N = self.N = lv.N
        self.nv = lv.n
        self.shell = shell
        self.params = shell_map_params(self.V, self.C) if shell else None
        # ---- edge DoFs: cell-local arrays in HyTeG's layout -> global edge ids
        self.edge_index = []  # per orientation: (points [m,3], first local index)
        off = 0
        for o in range(7):
            w = N if o < 6 else N - 1
            pts = np.array([(x, y, z) for z in range(w) for y in range(w - z) for x in range(w - z - y)], dtype=np.int64).reshape(-1, 3)
            self.edge_index.append((pts, off))
            off += tet_size(w)
        self.cell_edge_size = off
        key_to_edge = {}
        ends = []
        self.cell_edge_map = []
        for c in range(len(self.C)):
            vm = lv.cell_map[c]
            ids = np.empty(self.cell_edge_size, dtype=np.int64)
            for o in range(7):
                pts, first = self.edge_index[o]
                if len(pts) == 0:
                    continue
                a = pts + EDGE_ENDS[o, 0]
                b = pts + EDGE_ENDS[o, 1]
                ga = vm[po.cell_index(N, a[:, 0], a[:, 1], a[:, 2])]
                gb = vm[po.cell_index(N, b[:, 0], b[:, 1], b[:, 2])]
                for i, (p, q) in enumerate(zip(ga, gb)):
                    key = (min(p, q), max(p, q))
                    e = key_to_edge.get(key)
                    if e is None:
                        e = len(key_to_edge)
                        key_to_edge[key] = e
                        ends.append(key)
                    ids[first + i] = e
            self.cell_edge_map.append(ids)
        self.ne = len(key_to_edge)
        self.n = self.nv + self.ne
        self.edge_ends = np.array(ends, dtype=np.int64).reshape(-1, 2)
        self.key_to_edge = key_to_edge
        # computational coordinates and DoF types
        self.xyz = np.concatenate([lv.xyz, 0.5 * (lv.xyz[self.edge_ends[:, 0]] + lv.xyz[self.edge_ends[:, 1]])])
        et = np.empty(self.ne, dtype=np.int64)
        for e, (p, q) in enumerate(self.edge_ends):
            prim = tuple(sorted(set(lv.prim_of[p]) | set(lv.prim_of[q])))
            et[e] = self.p1.bc_types[self.p1.prim_flag[prim]]
        self.doftype = np.concatenate([lv.doftype, et])
        self._elements()

    def mask(self, flag):
        return (self.doftype & flag) != 0

    # -- micro elements: global DoF ids [ne,10], computational vertex coordinates [ne,4,3], macro-cell [ne]
    def _elements(self):
        lv, N = self.lv, self.N
        dofs, X, cellid = [], [], []
        anchors = np.array([(x, y, z) for z in range(N) for y in range(N - z) for x in range(N - z - y)], dtype=np.int64).reshape(-1, 3)
        for c, gv in enumerate(self.C):
            v = self.V[gv]
            vm = lv.cell_map[c]
            for t in range(6):
                a = anchors[anchors.sum(1) <= N - CELL_TYPE_SLACK[t]]
                if len(a) == 0:
                    continue
                vert = a[:, None, :] + CELL_TYPES[t][None, :, :]
                gid = vm[po.cell_index(N, vert[..., 0], vert[..., 1], vert[..., 2])]
                ed = np.empty((len(a), 6), dtype=np.int64)
                for k, (p, q) in enumerate(EDGE_PAIRS):
                    lo, hi = np.minimum(gid[:, p], gid[:, q]), np.maximum(gid[:, p], gid[:, q])
                    ed[:, k] = [self.key_to_edge[(int(i), int(j))] for i, j in zip(lo, hi)]
                dofs.append(np.concatenate([gid, self.nv + ed], axis=1))
                X.append(v[0] + (vert[..., 0:1] * (v[1] - v[0]) + vert[..., 1:2] * (v[2] - v[0]) + vert[..., 2:3] * (v[3] - v[0])) / N)
                cellid.append(np.full(len(a), c))
        self.el_dofs = np.concatenate(dofs)
        self.el_X = np.concatenate(X)
        self.el_cell = np.concatenate(cellid)

    def physical(self, x_comp, cells):
        return shell_eval_f(self.params[cells], x_comp) if self.shell else x_comp

    def dof_physical_coordinates(self):
        """position at which P2Function::interpolate evaluates its expression for every DoF (blended when the shell map is on; a DoF
        on a macro face / edge uses that primitive's map = the map of its first neighbour cell -- the same point up to rounding)"""
        if not self.shell:
            return self.xyz
        out = np.empty_like(self.xyz)
        for c in range(len(self.C)):
            vid = self.lv.cell_map[c]
            out[vid] = shell_eval_f(self.params[c], self.xyz[vid])
            eid = self.nv + self.cell_edge_map[c]
            out[eid] = shell_eval_f(self.params[c], self.xyz[eid])
        return out

    def element_matrices(self, form="divkgrad", k=coefficient_k, kvec=None, comp=(0, 0)):
        """[ne,10,10]; k: callable on physical points, or kvec: P2 DoF vector of the coefficient (evaluated with the P2 basis)"""
        rule = (VR1_POINTS, VR1_WEIGHTS) if form == "epsilon" else (XG4_POINTS, XG4_WEIGHTS)
        X = self.el_X
        J = np.stack([X[:, 1] - X[:, 0], X[:, 2] - X[:, 0], X[:, 3] - X[:, 0]], axis=2)  # columns = edge vectors
        det_aff = np.abs(np.linalg.det(J))
        A = np.zeros((len(X), 10, 10))
        P = self.params[self.el_cell] if self.shell else None
        for lam, w in zip(*rule):
            phi, gref = p2_shape(lam)
            xq = np.einsum("k,eki->ei", lam, X)
            if self.shell:
                DF = shell_eval_df(P, xq)
                M = DF @ J
                det_df = np.abs(np.linalg.det(DF))
                xp = shell_eval_f(P, xq)
            else:
                M = J
                det_df = 1.0
                xp = xq
            G = gref[None, :, :] @ np.linalg.inv(M)  # physical gradients [e,10,3]
            if form == "laplace":
                kq = 1.0
            elif kvec is not None:
                kq = kvec[self.el_dofs] @ phi
            else:
                kq = k(xp)
            s = w * kq * det_aff * det_df
            if form == "epsilon":
                c, d = comp  # test component c, trial component d: 2 mu eps(phi_j e_d) : eps(phi_i e_c)
                B = G[:, :, d, None] * G[:, None, :, c]  # d_d phi_i * d_c phi_j
                if c == d:
                    B = B + np.einsum("eik,ejk->eij", G, G)
                A += s[:, None, None] * B
            else:
                A += (s[:, None, None] if np.ndim(s) else s) * np.einsum("eik,ejk->eij", G, G)
        return A

    def div_element_matrices(self, comp):
        """[ne,4,10]: P2 -> P1, b_ij = - int psi_i d_comp phi_j  (forms/form_hyteg_generated/p2_to_p1/p2_to_p1_div_blending_q2.cpp);
        the transpose [ne,10,4] is p1_to_p2_divt_blending_q2 (P1 -> P2)"""
        X = self.el_X
        J = np.stack([X[:, 1] - X[:, 0], X[:, 2] - X[:, 0], X[:, 3] - X[:, 0]], axis=2)
        B = np.zeros((len(X), 4, 10))
        P = self.params[self.el_cell] if self.shell else None
        for lam, w in zip(VR1_POINTS, VR1_WEIGHTS):
            _, gref = p2_shape(lam)
            xq = np.einsum("k,eki->ei", lam, X)
            M = shell_eval_df(P, xq) @ J if self.shell else J
            G = gref[None, :, :] @ np.linalg.inv(M)
            B -= (w * np.abs(np.linalg.det(M)))[:, None, None] * lam[None, :, None] * G[:, None, :, comp]
        return B

    def div_matrix(self, comp):
        """P2ToP1ElementwiseBlendingDiv{x,y,z}Operator as a matrix [nv, n]"""
        B = self.div_element_matrices(comp)
        rows = np.repeat(self.el_dofs[:, :4], 10, axis=1).ravel()
        cols = np.tile(self.el_dofs, (1, 4)).ravel()
        M = sp.csr_matrix((B.ravel(), (rows, cols)), shape=(self.nv, self.n))
        M.sum_duplicates()
        return M

    def matrix(self, **kw):
        A = self.element_matrices(**kw)
        rows = np.repeat(self.el_dofs, 10, axis=1).ravel()
        cols = np.tile(self.el_dofs, (1, 10)).ravel()
        M = sp.csr_matrix((A.ravel(), (rows, cols)), shape=(self.n, self.n))
        M.sum_duplicates()
        return M

    # -- layout conversion (HyTeG macro-cell arrays <-> global vector)
    def to_cells(self, x):
        v = np.concatenate([x[self.lv.cell_map[c]] for c in range(len(self.C))])
        e = np.concatenate([x[self.nv + self.cell_edge_map[c]] for c in range(len(self.C))])
        return v, e

    def from_cells(self, v, e, tol=0.0):
        x = np.full(self.n, np.nan)
        mv, me = len(self.lv.pts), self.cell_edge_size
        for c in range(len(self.C)):
            for ids, seg in ((self.lv.cell_map[c], v[c * mv:(c + 1) * mv]), (self.nv + self.cell_edge_map[c], e[c * me:(c + 1) * me])):
                old = x[ids]
                bad = ~np.isnan(old) & (np.abs(old - seg) > tol)
                if bad.any():
                    raise AssertionError("copies of a DoF disagree in macro-cell %d: %g" % (c, np.abs(old - seg)[bad].max()))
                x[ids] = seg
        return x


def apply(o, M, x, dst, flag, add=False):
    """P2ElementwiseOperator::apply: DoFs of primitives whose boundary type is not selected stay untouched"""
    m = o.mask(flag)
    y = M @ x
    out = dst.copy()
    out[m] = (out[m] + y[m]) if add else y[m]
    return out


# ---- quadratic grid transfer ---------------------------------------------------------------------------------------------------------

def quadratic_prolongation(coarse, fine, tol=1e-9):
    """P2toP2QuadraticProlongation as a matrix [fine.n, coarse.n]: nodal interpolation -- the coarse P2 function evaluated at every fine
    DoF position (reference src/hyteg/gridtransferoperators/P2toP2QuadraticProlongation.cpp, generated prolongate_3D_macrocell_P2_push_*).
    P2toP2QuadraticRestriction (P2toP2QuadraticRestriction.cpp:131-251) is its transpose."""
    from scipy.spatial import cKDTree
    lat = np.array([(i, j, k, 4 - i - j - k) for i in range(5) for j in range(5 - i) for k in range(5 - i - j)], dtype=np.float64) / 4.0
    shapes = np.array([p2_shape(l)[0] for l in lat])  # [35,10]
    tree = cKDTree(fine.xyz)
    X = np.einsum("pk,eki->epi", lat, coarse.el_X)  # [ne,35,3]
    d, idx = tree.query(X.reshape(-1, 3))
    assert d.max() < tol, d.max()
    rows = np.repeat(idx, 10)
    cols = np.repeat(coarse.el_dofs, 35, axis=0).ravel()
    vals = np.tile(shapes, (len(coarse.el_X), 1)).ravel()
    # every (fine, coarse) pair appears once per coarse element containing the fine point, always with the same value: keep one
    key = rows * coarse.n + cols
    _, first = np.unique(key, return_index=True)
    P = sp.csr_matrix((vals[first], (rows[first], cols[first])), shape=(fine.n, coarse.n))
    P.eliminate_zeros()
    return P


# ---- smoothers and multigrid on assembled P2 operators (same algorithms as oracle/p1_oracle.py, on matrices) ----------------------------

SMOOTH_FLAG = po.SMOOTH_FLAG


def estimate_radius(M, x, iters):
    """chebyshev::estimateRadius: power iteration on D^-1 A, flag All (numerictools/SpectrumEstimation.hpp:56-85)"""
    d = M.diagonal()
    op = lambda v: (M @ v) / d
    x = x / np.sqrt(x @ x)
    aux = op(x)
    radius = 0.0
    for _ in range(iters):
        x = aux / np.sqrt(aux @ aux)
        aux = op(x)
        radius = x @ aux
    return radius


def chebyshev(o, M, x, b, coeffs, flag=SMOOTH_FLAG):
    """ChebyshevSmoother::solve (solvers/ChebyshevSmoother.hpp:179-222) with the inverse diagonal as preconditioner"""
    m = o.mask(flag)
    d = M.diagonal()
    x = x.copy()
    tmp1 = np.zeros(o.n)
    tmp1[m] = ((b - M @ x) / d)[m]
    x[m] += coeffs[0] * tmp1[m]
    for k in range(1, len(coeffs)):
        t = np.zeros(o.n)
        t[m] = ((M @ tmp1) / d)[m]
        tmp1 = t
        x[m] += coeffs[k] * tmp1[m]
    return x


def cg(o, M, x, b, max_iter, rel_tol, abs_tol, flag=SMOOTH_FLAG):
    m = o.mask(flag)
    x = x.copy()
    r = np.zeros(o.n)
    r[m] = (b - M @ x)[m]
    p = r.copy()
    rs = r @ r
    r0 = np.sqrt(rs)
    if r0 < abs_tol:
        return x
    for _ in range(max_iter):
        ap = np.zeros(o.n)
        ap[m] = (M @ p)[m]
        alpha = rs / (p @ ap)
        x[m] += alpha * p[m]
        r[m] -= alpha * ap[m]
        rn = r @ r
        if np.sqrt(rn) / r0 < rel_tol or np.sqrt(rn) < abs_tol:
            break
        p = r + (rn / rs) * p
        rs = rn
    return x


def vcycle(os_, Ms, Ps, coeffs, x, b, L, min_level, pre, post, flag=SMOOTH_FLAG):
    """GeometricMultigridSolver::solveRecursively (solvers/GeometricMultigridSolver.hpp:207-312) with the Chebyshev smoother,
    quadratic transfer (Ps[l]: level l-1 -> l) and CG (1e-14, at most 1000 iterations) on min_level; dicts indexed by level"""
    o, M = os_[L], Ms[L]
    if L == min_level:
        x[L] = cg(o, M, x[L], b[L], 1000, 1e-14, 1e-14)
        return
    for _ in range(pre):
        x[L] = chebyshev(o, M, x[L], b[L], coeffs)
    m = o.mask(flag)
    tmp = np.zeros(o.n)
    tmp[m] = (b[L] - M @ x[L])[m]
    mc = os_[L - 1].mask(flag)
    rc_ = b[L - 1].copy()
    rc_[mc] = (Ps[L].T @ tmp)[mc]
    b[L - 1] = rc_
    x[L - 1] = np.zeros(os_[L - 1].n)
    vcycle(os_, Ms, Ps, coeffs, x, b, L - 1, min_level, pre, post, flag)
    x[L][m] += (Ps[L] @ x[L - 1])[m]
    for _ in range(post):
        x[L] = chebyshev(o, M, x[L], b[L], coeffs)
