"""CPU oracle for the P1 constant-stencil multigrid path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may import
this module; the product (hyteg_b200/) never does.

What it restates (reference = HyTeG @ a7ecfdf6, paths relative to the reference tree):
  * micro mesh of a macro-tet: the six micro-cell types and their vertex offsets,
      src/hyteg/volumedofspace/CellDoFIndexing.hpp:155-198 (getMicroVerticesFromMicroCell)
  * the operator: a constant-stencil P1 operator is the sum of the local stiffness matrices of the
      micro-cells around a micro-vertex (P1Elements.hpp:465-515 for macro-cells, :626-1017 for the
      macro-face/edge/vertex stencils which add up the neighbouring macro-cells' parts); that is the
      assembled P1 stiffness matrix of the refined mesh, which is what this oracle builds
      (local matrix: forms/form_fenics_generated/p1_tet_diffusion.h = |T| grad phi_i . grad phi_j)
  * DoFType selection at macro-primitive granularity, P1Operator.hpp:211-315, types.hpp:29-44
  * smooth_jac: P1Operator.hpp:429-447; weighted Jacobi driver: WeightedJacobiSmoother.hpp:46-65
  * smooth_sor ordering: P1Operator.hpp:348-418 + sor_3D_macrocell_P1.cpp:53-76 (vertices, then
      edges, then faces, then cells; lexicographic inside a primitive; primitives of one class do
      not see each other's updates)
  * Chebyshev: ChebyshevSmoother.hpp:179-222; power iteration: SpectrumEstimation.hpp:56-85
  * CG: CGSolver.hpp:94-207 ; V-cycle: GeometricMultigridSolver.hpp:207-312
  * transfer: linear interpolation / its transpose, P1toP1LinearProlongation.cpp:194-410,
      P1toP1LinearRestriction.cpp:169-346

Pinning (see tests/test_oracle_pinning.py and oracle/README.md): the oracle is checked against the
reference's own known-answer and golden values -- stencil symmetry and row sums
(VertexDoFStencilAssemblyTest.cpp:64-238), A*const = A*linear = 0 < 2.8e-13
(P1LaplaceOperator3DTest.cpp), the V(3,3)+GS convergence bound 3.2e-2 of P1GMG3DConvergenceTest.cpp:148,
and against the reference's generated macro-cell kernels compiled from the reference tree
(oracle/_ref, built by oracle/Makefile).
"""
import numpy as np
import scipy.sparse as sp

INNER, DIRICHLET, NEUMANN, FREESLIP, ALL = 1, 2, 4, 8, 15

MICRO_CELLS = np.array([
    [[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]],  # WHITE_UP
    [[1, 0, 0], [1, 1, 0], [0, 1, 0], [1, 0, 1]],  # BLUE_UP
    [[1, 0, 0], [0, 1, 0], [1, 0, 1], [0, 0, 1]],  # GREEN_UP
    [[1, 1, 0], [1, 1, 1], [0, 1, 1], [1, 0, 1]],  # WHITE_DOWN
    [[1, 0, 1], [0, 1, 1], [0, 0, 1], [0, 1, 0]],  # BLUE_DOWN
    [[0, 1, 0], [1, 1, 0], [1, 0, 1], [0, 1, 1]],  # GREEN_DOWN
])


def cell_index(N, x, y, z):
    """linear index in the macro-cell array, MacroCellIndexing.hpp:44-52 (width = N+1)"""
    w = N + 1
    wz = w - z
    tet = lambda k: k * (k + 1) * (k + 2) // 6
    return tet(w) - tet(wz) + y * (wz + 1) - (y * (y + 1)) // 2 + x


def cell_points(N):
    """all (x,y,z) of a macro-cell array in storage order"""
    pts = [(x, y, z) for z in range(N + 1) for y in range(N + 1 - z) for x in range(N + 1 - z - y)]
    return np.array(pts, dtype=np.int64)


def p1_element_matrices(P):
    """P: [n,4,3] vertex coordinates -> [n,4,4] P1 stiffness matrices"""
    J = np.stack([P[:, 1] - P[:, 0], P[:, 2] - P[:, 0], P[:, 3] - P[:, 0]], axis=2)  # columns = edges
    det = np.linalg.det(J)
    Jinv = np.linalg.inv(J)  # rows = grad lambda_1..3
    G = np.concatenate([-Jinv.sum(axis=1, keepdims=True), Jinv], axis=1)  # [n,4,3]
    vol = np.abs(det) / 6.0
    return vol[:, None, None] * np.einsum("nik,njk->nij", G, G)


class Level:
    pass


class P1Oracle:
    """Global (unique-DoF) view of a P1 function space on a refined tetrahedral macro mesh."""

    def __init__(self, vertices, cells, min_level, max_level, flag_on_boundary=1, flag_inner=0,
                 bc_types=(INNER, DIRICHLET, NEUMANN, FREESLIP), primitive_flags=None):
        self.V = np.asarray(vertices, dtype=np.float64).reshape(-1, 3)
        self.C = np.asarray(cells, dtype=np.int64).reshape(-1, 4)
        self.min_level, self.max_level = min_level, max_level
        self.bc_types = bc_types
        self._boundary_flags(flag_on_boundary, flag_inner, primitive_flags)
        self.levels = {}
        for L in range(min_level, max_level + 1):
            self.levels[L] = self._build_level(L)
        for L in range(min_level + 1, max_level + 1):
            self.levels[L].P = self._build_prolongation(L)

    # -- macro topology -----------------------------------------------------------------------
    def _boundary_flags(self, fb, fi, primitive_flags):
        """SetupPrimitiveStorage::onBoundary / setMeshBoundaryFlagsOnBoundary(fb, fi, true)"""
        from itertools import combinations
        face_count = {}
        for c in self.C:
            for f in combinations(sorted(c), 3):
                face_count[f] = face_count.get(f, 0) + 1
        on_bnd = set()
        for f, n in face_count.items():
            if n == 1:
                for k in (1, 2, 3):
                    for s in combinations(f, k):
                        on_bnd.add(s)
        self.prim_flag = {}
        for c in self.C:
            sc = tuple(sorted(c))
            self.prim_flag[sc] = fi  # highestDimensionAlwaysInner
            for k in (1, 2, 3):
                for s in combinations(sc, k):
                    self.prim_flag[s] = fb if s in on_bnd else fi
        if primitive_flags:
            self.prim_flag.update(primitive_flags)

    # -- one refinement level -----------------------------------------------------------------
    def _build_level(self, L):
        lv = Level()
        N = 1 << L
        lv.N = N
        pts = cell_points(N)
        lv.pts = pts
        key_to_id = {}
        lv.cell_map = []
        prim_of = []
        for c, gv in enumerate(self.C):
            ids = np.empty(len(pts), dtype=np.int64)
            for i, (x, y, z) in enumerate(pts):
                b = (N - x - y - z, x, y, z)
                key = tuple(sorted((int(gv[j]), int(b[j])) for j in range(4) if b[j] > 0))
                g = key_to_id.get(key)
                if g is None:
                    g = len(key_to_id)
                    key_to_id[key] = g
                    prim_of.append(tuple(k[0] for k in key))
                ids[i] = g
            lv.cell_map.append(ids)
        lv.n = len(key_to_id)
        lv.key_to_id = key_to_id
        lv.prim_of = prim_of
        lv.dim = np.array([len(p) - 1 for p in prim_of])
        lv.doftype = np.array([self.bc_types[self.prim_flag[p]] for p in prim_of])
        # coordinates
        lv.xyz = np.zeros((lv.n, 3))
        for c, gv in enumerate(self.C):
            v = self.V[gv]
            X = v[0] + (pts[:, 0:1] * (v[1] - v[0]) + pts[:, 1:2] * (v[2] - v[0]) + pts[:, 2:3] * (v[3] - v[0])) / N
            lv.xyz[lv.cell_map[c]] = X
        # micro elements + assembly
        rows, cols, vals = [], [], []
        lv.elements = []
        anchors = np.array([(i, j, k) for k in range(N) for j in range(N) for i in range(N)], dtype=np.int64)
        for c, gv in enumerate(self.C):
            v = self.V[gv]
            for t in range(6):
                vert = anchors[:, None, :] + MICRO_CELLS[t][None, :, :]  # [n,4,3]
                inside = (vert.sum(axis=2) <= N).all(axis=1)
                vert = vert[inside]
                if len(vert) == 0:
                    continue
                lin = cell_index(N, vert[..., 0], vert[..., 1], vert[..., 2])
                gid = lv.cell_map[c][lin]  # [n,4]
                X = v[0] + (vert[..., 0:1] * (v[1] - v[0]) + vert[..., 1:2] * (v[2] - v[0]) + vert[..., 2:3] * (v[3] - v[0])) / N
                K = p1_element_matrices(X)
                rows.append(np.repeat(gid, 4, axis=1).ravel())
                cols.append(np.tile(gid, (1, 4)).ravel())
                vals.append(K.ravel())
                lv.elements.append(gid)
        lv.elements = np.concatenate(lv.elements)
        lv.A = sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(lv.n, lv.n))
        lv.A.sum_duplicates()
        lv.diag = lv.A.diagonal()
        return lv

    def _build_prolongation(self, L):
        """linear interpolation level L-1 -> L: coarse points injected, edge midpoints averaged"""
        fine, coarse = self.levels[L], self.levels[L - 1]
        rows, cols, vals = [], [], []
        done = set()
        for key, g in coarse.key_to_id.items():
            fkey = tuple((v, 2 * b) for v, b in key)
            rows.append(fine.key_to_id[fkey]); cols.append(g); vals.append(1.0)
        inv_key = {g: key for key, g in coarse.key_to_id.items()}
        edges = set()
        for el in coarse.elements:
            for a in range(4):
                for b in range(a + 1, 4):
                    e = (min(el[a], el[b]), max(el[a], el[b]))
                    edges.add(e)
        for (a, b) in edges:
            ka, kb = dict(inv_key[a]), dict(inv_key[b])
            mid = {}
            for v, w in ka.items():
                mid[v] = mid.get(v, 0) + w
            for v, w in kb.items():
                mid[v] = mid.get(v, 0) + w
            fkey = tuple(sorted(mid.items()))
            f = fine.key_to_id[fkey]
            rows += [f, f]; cols += [a, b]; vals += [0.5, 0.5]
        return sp.csr_matrix((vals, (rows, cols)), shape=(fine.n, coarse.n))

    # -- layout conversion ----------------------------------------------------------------------
    def to_cells(self, x, L, cells=None):
        """global vector -> concatenated per-macro-cell arrays (HyTeG layout)"""
        lv = self.levels[L]
        cells = range(len(self.C)) if cells is None else cells
        return np.concatenate([x[lv.cell_map[c]] for c in cells])

    def from_cells(self, arr, L, cells=None, check=True, tol=0.0):
        """per-macro-cell arrays -> global vector; verifies that all copies of a DoF agree"""
        lv = self.levels[L]
        cells = list(range(len(self.C))) if cells is None else list(cells)
        m = len(lv.pts)
        x = np.full(lv.n, np.nan)
        for i, c in enumerate(cells):
            seg = arr[i * m:(i + 1) * m]
            ids = lv.cell_map[c]
            if check:
                old = x[ids]
                bad = ~np.isnan(old) & (np.abs(old - seg) > tol)
                if bad.any():
                    raise AssertionError(f"copies of a DoF disagree in macro-cell {c}: max diff {np.abs(old - seg)[bad].max()}")
            x[ids] = seg
        return x

    def mask(self, L, flag):
        return (self.levels[L].doftype & flag) != 0


# ================================================================================================
#  Algorithms (global vectors; "flag" selects DoFs by the DoFType of their macro primitive)
# ================================================================================================

SMOOTH_FLAG = INNER | NEUMANN | FREESLIP


def apply(o, x, dst, L, flag, add=False):
    """P1Operator::apply (P1Operator.hpp:192-320): rows of excluded primitives are left untouched"""
    lv = o.levels[L]
    m = o.mask(L, flag)
    y = lv.A @ x
    out = dst.copy()
    out[m] = (out[m] + y[m]) if add else y[m]
    return out


def smooth_jac(o, rhs, src, omega, L, flag, dst=None):
    """P1Operator::smooth_jac (P1Operator.hpp:429-447): dst = src + omega D^-1 (rhs - A src)"""
    lv = o.levels[L]
    m = o.mask(L, flag)
    out = src.copy() if dst is None else dst.copy()
    r = rhs - lv.A @ src
    out[m] = src[m] + omega * (r[m] / lv.diag[m])
    return out


def _lex_order(o, L):
    """HyTeG's sweep order: class by class (vertices, edges, faces, cells), lexicographic inside every primitive in the
    primitive's OWN index space: a macro-cell in the mesh's vertex order (x fastest, sor_3D_macrocell_P1.cpp:53-76); a
    macro-face / macro-edge in the order of its vertices SORTED by mesh vertex id (MeshInfo::addFace / addEdge sort them,
    mesh/MeshInfo.cpp:37-72, and SetupPrimitiveStorage.cpp:126-271 keeps that order), column index fastest
    (sor_3D_macroface_P1*.cpp, P1Operator.hpp smooth_sor_edge).  Pinned against the reference itself
    (tests/test_oracle_pinning.py, tests/golden/ref_p1.npz)."""
    lv = o.levels[L]
    if hasattr(lv, "lex"):
        return lv.lex
    seen = np.zeros(lv.n, dtype=bool)
    order = [[] for _ in range(4)]
    for c in range(len(o.C)):
        for g in lv.cell_map[c]:
            if not seen[g]:
                seen[g] = True
                order[lv.dim[g]].append(g)
    inv_key = {g: key for key, g in lv.key_to_id.items()}
    # edges: position = weight of the higher vertex; faces: (row = weight of the highest vertex, column = weight of the middle one)
    order[1].sort(key=lambda g: (lv.prim_of[g], inv_key[g][1][1]))
    order[2].sort(key=lambda g: (lv.prim_of[g], inv_key[g][2][1], inv_key[g][1][1]))
    lv.lex = order
    return order


def smooth_sor(o, x, rhs, omega, L, flag, backwards=False):
    """P1Operator::smooth_sor (P1Operator.hpp:348-418) in the reference's hybrid ordering: block Gauss-Seidel over the classes
    vertices -> edges -> faces -> cells (reversed when backwards), block Jacobi between the primitives of a class,
    lexicographic inside a primitive.

    One more property of the reference is restated, found by pinning against the reference itself: a macro-face reads the
    points of its neighbour cells through its ghost layers, which are refreshed only by the communicate<Cell,Face> in FRONT of
    the sweep (P1Operator.hpp:373-379); the communicate<Edge,Face> between the edge and face classes (:403) refreshes the
    face's own boundary only.  In a forward sweep a face-interior point therefore sees the PRE-SWEEP value of a neighbour that
    lies on a macro-edge of a neighbour cell which is not an edge of the face itself (stencil direction through the ghost layer).
    The mirror image in a backward sweep: a macro-edge point sees the PRE-SWEEP value of a neighbour on a macro-face that does
    not contain the edge (reached through the edge's cell ghost layer, which travels Cell -> Face -> Edge and so carries the
    cell's copy of that face from before the face class ran)."""
    lv = o.levels[L]
    A = lv.A
    indptr, indices, data = A.indptr, A.indices, A.data
    m = o.mask(L, flag)
    x0 = x
    x = x.copy()
    order = _lex_order(o, L)
    prim = lv.prim_of
    classes = [3, 2, 1, 0] if backwards else [0, 1, 2, 3]
    for d in classes:
        snap = x.copy()  # primitives of the same class do not see each other's updates
        pts = order[d][::-1] if backwards else order[d]
        for g in pts:
            if not m[g]:
                continue
            s = rhs[g]
            dg = 0.0
            pg = set(prim[g])
            for k in range(indptr[g], indptr[g + 1]):
                j = indices[k]
                if j == g:
                    dg = data[k]
                elif lv.dim[j] == d and prim[j] != prim[g]:
                    s -= data[k] * snap[j]
                elif d == 2 and not backwards and lv.dim[j] < 2 and not set(prim[j]) <= pg:
                    s -= data[k] * x0[j]  # stale ghost layer, see above
                elif d == 1 and backwards and lv.dim[j] == 2 and not pg <= set(prim[j]):
                    s -= data[k] * x0[j]  # stale ghost layer, backward sweep
                else:
                    s -= data[k] * x[j]
            x[g] = (1.0 - omega) * x[g] + omega * s / dg
    return x


def _multicolour_phases(o, L, interior):
    """Phases of the DEVICE ordering (hyteg_b200/csrc/runtime_operator.inc hb_operator_smooth_sor_ordered), each a list of
    index arrays that are updated simultaneously, in forward order:
      0      macro vertices
      1, 2   macro-edge points by parity of the barycentric weight of the edge's higher vertex
      3..5   macro-face points by (w_mid + 2 w_high) mod 3 (weights w.r.t. the face's vertices sorted by global id)
      then the macro-cell interiors in four colours (x + 2y + 3z) mod 4 (a proper colouring of the 15-point stencil graph)."""
    lv = o.levels[L]
    key = ("mc", interior)
    if hasattr(lv, "mc") and key in lv.mc:
        return lv.mc[key]
    inv_key = {g: k for k, g in lv.key_to_id.items()}
    ph = [[] for _ in range(6)]
    for g in range(lv.n):
        k = inv_key[g]
        if len(k) == 1:
            ph[0].append(g)
        elif len(k) == 2:
            ph[1 + (k[1][1] & 1)].append(g)
        elif len(k) == 3:
            ph[3 + (k[1][1] + 2 * k[2][1]) % 3].append(g)
    phases = [np.array(p, dtype=np.int64) for p in ph]
    pts, N = lv.pts, lv.N
    inter = (pts[:, 0] > 0) & (pts[:, 1] > 0) & (pts[:, 2] > 0) & (pts.sum(1) < N)
    idx = np.nonzero(inter)[0]
    x_, y_, z_ = pts[idx, 0], pts[idx, 1], pts[idx, 2]
    assert interior == "c4"
    col = (x_ + 2 * y_ + 3 * z_) % 4
    for c in range(4):
        phases.append(np.concatenate([lv.cell_map[k][idx[col == c]] for k in range(len(o.C))]) if len(idx) else np.zeros(0, dtype=np.int64))
    if not hasattr(lv, "mc"):
        lv.mc = {}
    lv.mc[key] = phases
    return phases


def smooth_sor_multicolour(o, x, rhs, omega, L, flag, backwards=False, interior="c4"):
    """The device's multicolour restatement of P1Operator::smooth_sor: same class order as the reference (vertices, edges,
    faces, cells; reversed when backwards) and block Jacobi between the primitives of a class, but colours instead of the
    lexicographic order inside a primitive.  Every phase is a simultaneous update."""
    lv = o.levels[L]
    m = o.mask(L, flag)
    x = x.copy()
    phases = _multicolour_phases(o, L, interior)
    for g in (phases[::-1] if backwards else phases):
        g = g[m[g]]
        if len(g) == 0:
            continue
        r = rhs[g] - lv.A[g] @ x
        x[g] = x[g] + omega * r / lv.diag[g]
    return x


def cheb_coefficients(order, lower, upper):
    """ChebyshevSmoother::setupCoefficientsInternal (ChebyshevSmoother.hpp:329-539).  The reference tabulates closed forms for
    orders 1..10; they are the monomial coefficients c_i of the polynomial p with
        1 - lambda p(lambda) = T_k( (theta - lambda) / delta ) / T_k( theta / delta ),   p = sum_i c_i lambda^i,
    (the smoother applies x += sum_i c_i (D^-1 A)^i D^-1 r, ChebyshevSmoother.hpp:179-222).  Restated here from that
    definition with exact polynomial algebra; pinned against the reference's own tables for orders 1..10
    (tests/golden/ref_p1.npz: cheb_coeffs_order*)."""
    from numpy.polynomial import chebyshev as C, polynomial as P
    theta = 0.5 * (upper + lower)
    delta = 0.5 * (upper - lower)
    tk = C.cheb2poly([0.0] * order + [1.0])  # T_k(s) in powers of s
    # substitute s = (theta - lambda) / delta
    lin = np.array([theta / delta, -1.0 / delta])
    q = np.zeros(1)
    pw = np.ones(1)
    for a in tk:
        q = P.polyadd(q, a * pw)
        pw = P.polymul(pw, lin)
    q = q / P.polyval(0.0, q)  # T_k((theta - lambda)/delta) / T_k(theta/delta): value 1 at lambda = 0
    one_minus = P.polysub(np.ones(1), q)  # = lambda p(lambda): constant term vanishes
    return [float(c) for c in one_minus[1:order + 1]]


def chebyshev(o, x, b, coeffs, L, flag=SMOOTH_FLAG):
    """ChebyshevSmoother::solve (ChebyshevSmoother.hpp:179-222); tmp functions are zero on excluded DoFs"""
    lv = o.levels[L]
    m = o.mask(L, flag)
    x = x.copy()
    tmp2 = np.zeros(lv.n)
    tmp1 = np.zeros(lv.n)
    tmp2[m] = (b - lv.A @ x)[m]
    tmp1[m] = tmp2[m] / lv.diag[m]
    x[m] += coeffs[0] * tmp1[m]
    for k in range(1, len(coeffs)):
        tmp2[m] = (lv.A @ tmp1)[m]
        tmp1[m] = tmp2[m] / lv.diag[m]
        x[m] += coeffs[k] * tmp1[m]
    return x


def estimate_radius(o, x, iters, L):
    """estimateSpectralRadiusWithPowerIteration on D^-1 A, flag All (SpectrumEstimation.hpp:56-85)"""
    lv = o.levels[L]
    op = lambda v: (lv.A @ v) / lv.diag
    x = x / np.sqrt(x @ x)
    aux = op(x)
    radius = 0.0
    for _ in range(iters):
        x = aux / np.sqrt(aux @ aux)
        aux = op(x)
        radius = x @ aux
    return radius


def cg(o, x, b, L, max_iter, rel_tol=1e-16, abs_tol=1e-16, flag=SMOOTH_FLAG):
    """CGSolver::solve with identity preconditioner (CGSolver.hpp:94-207)"""
    lv = o.levels[L]
    m = o.mask(L, flag)
    x = x.copy()
    r = np.zeros(lv.n)
    p = np.zeros(lv.n)
    r[m] = (b - lv.A @ x)[m]
    z = r.copy()
    p[m] = z[m]
    prsold = r[m] @ z[m]
    res0 = np.sqrt(r[m] @ r[m])
    if res0 < abs_tol:
        return x, 0
    it = max_iter
    for i in range(max_iter):
        ap = np.zeros(lv.n)
        ap[m] = (lv.A @ p)[m]
        pAp = p[m] @ ap[m]
        alpha = prsold / pAp
        x[m] += alpha * p[m]
        r[m] -= alpha * ap[m]
        rs = r[m] @ r[m]
        if np.sqrt(rs) / res0 < rel_tol or np.sqrt(rs) < abs_tol:
            it = i
            break
        z = r.copy()
        prsnew = r[m] @ z[m]
        beta = prsnew / prsold
        p[m] = z[m] + beta * p[m]
        prsold = prsnew
    return x, it


def restrict(o, f_fine, c_old, L, flag):
    """P1toP1LinearRestriction::restrict3D: coarse = P^T fine on the selected coarse DoFs"""
    m = o.mask(L - 1, flag)
    out = c_old.copy()
    out[m] = (o.levels[L].P.T @ f_fine)[m]
    return out


def prolongate_and_add(o, c, f_old, L_coarse, flag):
    """P1toP1LinearProlongation::prolongateAndAdd: fine += P coarse on the selected fine DoFs"""
    m = o.mask(L_coarse + 1, flag)
    out = f_old.copy()
    out[m] += (o.levels[L_coarse + 1].P @ c)[m]
    return out


def vcycle(o, x, b, L, min_level, smoother, coarse_solver, pre=3, post=3, flag=SMOOTH_FLAG):
    """GeometricMultigridSolver::solveRecursively (GeometricMultigridSolver.hpp:207-312).

    x, b: dicts level -> global vector (coarse levels of b and x are overwritten, as in the reference)
    smoother(o, x, b, L) -> x ; coarse_solver(o, x, b, L) -> x
    """
    if L == min_level:
        x[L] = coarse_solver(o, x[L], b[L], L)
        return
    for _ in range(pre):
        x[L] = smoother(o, x[L], b[L], L)
    lv = o.levels[L]
    m = o.mask(L, flag)
    tmp = np.zeros(lv.n)
    tmp[m] = (b[L] - lv.A @ x[L])[m]
    b[L - 1] = restrict(o, tmp, b[L - 1], L, flag)
    x[L - 1] = np.zeros(o.levels[L - 1].n)
    vcycle(o, x, b, L - 1, min_level, smoother, coarse_solver, pre, post, flag)
    x[L] = prolongate_and_add(o, x[L - 1], x[L], L - 1, flag)
    for _ in range(post):
        x[L] = smoother(o, x[L], b[L], L)
