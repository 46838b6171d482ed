"""Macro meshes (MeshInfo role): vertex coordinates + tetrahedra in mesh order.

cube_6el reproduces reference data/meshes/3D/cube_6el.msh; the spherical shell follows the
construction of MeshInfo::meshSphericalShell (reference src/hyteg/mesh/MeshGenSphericalShell.cpp):
an icosahedron's 10 diamonds, each tangentially subdivided and radially layered, every prism cut
into 3 tetrahedra (60*(ntan-1)^2*(nrad-1) cells).
"""
import json
import os

import numpy as np


def cube_6el():
    v = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], dtype=np.float64)
    c = np.array([[0, 1, 3, 4], [1, 3, 4, 5], [1, 2, 3, 5], [4, 5, 3, 7], [5, 6, 3, 7], [5, 2, 3, 6]], dtype=np.int32)
    return v, c


def tet_1el():
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], dtype=np.float64)
    c = np.array([[0, 1, 2, 3]], dtype=np.int32)
    return v, c


def from_gmsh(path):
    """Gmsh 2.2 ASCII reader for tetrahedral meshes (element type 4)."""
    with open(path) as f:
        tok = f.read().split("\n")
    i = tok.index("$Nodes")
    n = int(tok[i + 1])
    ids, v = {}, []
    for k in range(n):
        p = tok[i + 2 + k].split()
        ids[int(p[0])] = k
        v.append([float(p[1]), float(p[2]), float(p[3])])
    i = tok.index("$Elements")
    n = int(tok[i + 1])
    c = []
    for k in range(n):
        p = tok[i + 2 + k].split()
        if int(p[1]) != 4:
            continue
        ntags = int(p[2])
        c.append([ids[int(q)] for q in p[3 + ntags:3 + ntags + 4]])
    return np.array(v, dtype=np.float64), np.array(c, dtype=np.int32)


def golden(name):
    """Meshes stored in tests/golden/meshes.json (extracted from the reference's data/meshes/3D)."""
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "..", "tests", "golden", "meshes.json")) as f:
        d = json.load(f)[name]
    return np.array(d["vertices"], dtype=np.float64), np.array(d["cells"], dtype=np.int32)


def cuboid(nx, ny, nz, lo=(0, 0, 0), hi=(1, 1, 1)):
    """Structured cuboid of nx*ny*nz cubes, each split into 6 tetrahedra around the main diagonal."""
    xs = np.linspace(lo[0], hi[0], nx + 1)
    ys = np.linspace(lo[1], hi[1], ny + 1)
    zs = np.linspace(lo[2], hi[2], nz + 1)
    idx = lambda i, j, k: (k * (ny + 1) + j) * (nx + 1) + i
    v = np.array([[xs[i], ys[j], zs[k]] for k in range(nz + 1) for j in range(ny + 1) for i in range(nx + 1)])
    cells = []
    # Kuhn triangulation: 6 tets per cube along paths from corner 000 to 111
    perms = [(0, 1, 2), (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)]
    for k in range(nz):
        for j in range(ny):
            for i in range(nx):
                for p in perms:
                    cur = [i, j, k]
                    tet = [idx(*cur)]
                    for ax in p:
                        cur[ax] += 1
                        tet.append(idx(*cur))
                    cells.append(tet)
    return v, np.array(cells, dtype=np.int32)


def spherical_shell(ntan, nrad, rmin=0.5, rmax=1.0):
    """Icosahedral thick spherical shell with 60*(ntan-1)^2*(nrad-1) tetrahedra."""
    radii = np.linspace(rmin, rmax, nrad)
    # icosahedron vertices
    t = (1.0 + 5 ** 0.5) / 2.0
    ico = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                    [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    ico /= np.linalg.norm(ico[0])
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
             (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
             (8, 6, 7), (9, 8, 1)]
    n = ntan - 1
    vmap, verts = {}, []

    def vid(p, r):
        key = (tuple(np.round(p, 12)), r)
        if key not in vmap:
            vmap[key] = len(verts)
            verts.append(p * radii[r])
        return vmap[key]

    cells = []
    for (a, b, c) in faces:
        A, B, Cc = ico[a], ico[b], ico[c]

        def pt(i, j):
            q = (A * (n - i - j) + B * i + Cc * j) / n
            return q / np.linalg.norm(q)

        tris = []
        for i in range(n):
            for j in range(n - i):
                tris.append(((i, j), (i + 1, j), (i, j + 1)))
                if i + j < n - 1:
                    tris.append(((i + 1, j), (i + 1, j + 1), (i, j + 1)))
        for r in range(nrad - 1):
            for tri in tris:
                lo = [vid(pt(*q), r) for q in tri]
                hi = [vid(pt(*q), r + 1) for q in tri]
                # consistent prism split: order by global vertex id of the lower triangle
                order = np.argsort(lo)
                l = [lo[o] for o in order]
                h = [hi[o] for o in order]
                cells.append([l[0], l[1], l[2], h[2]])
                cells.append([l[0], l[1], h[1], h[2]])
                cells.append([l[0], h[0], h[1], h[2]])
    return np.array(verts), np.array(cells, dtype=np.int32)


def partition_radial_layers(vertices, cells, n_parts):
    """Equally sized blocks of macro-cells by distance from the origin: on meshSphericalShell( ntan, n_parts + 1 ) block r is the r-th
    radial layer, so that a rank talks to at most two neighbours through the two spherical surfaces of its layer.  (What a user gets
    from SetupPrimitiveStorage::setTargetRank with the layer index; measured against partition_greedy on shell( 2, 9 ) over 8 ranks:
    half the ghost volume, 2 instead of up to 7 neighbour ranks.)  Returns the target rank of every cell."""
    v, c = np.asarray(vertices, dtype=float).reshape(-1, 3), np.asarray(cells).reshape(-1, 4)
    n = len(c)
    if n_parts <= 1:
        return np.zeros(n, dtype=np.int32)
    order = np.argsort(np.linalg.norm(v[c].mean(axis=1), axis=1), kind="stable")
    rank = np.zeros(n, dtype=np.int32)
    for r in range(n_parts):
        rank[order[r * n // n_parts:(r + 1) * n // n_parts]] = r
    return rank


def partition_greedy(cells, n_parts):
    """Connected, equally sized blocks of macro-cells: n_parts regions grown breadth-first over the face-adjacency graph from seeds
    spread over the mesh, one cell per region per round -- in the spirit of the reference's loadbalancing::greedy
    (src/hyteg/primitivestorage/loadbalancing/SimpleBalancer.hpp:56-78).  Returns the target rank of every cell."""
    from collections import deque
    cells = np.asarray(cells)
    n = len(cells)
    if n_parts <= 1:
        return np.zeros(n, dtype=np.int32)
    face_cells = {}
    for c, t in enumerate(cells):
        for f in ((0, 1, 2), (0, 1, 3), (0, 2, 3), (1, 2, 3)):
            face_cells.setdefault(tuple(sorted(int(t[i]) for i in f)), []).append(c)
    adj = [[] for _ in range(n)]
    for cs in face_cells.values():
        if len(cs) == 2:
            adj[cs[0]].append(cs[1])
            adj[cs[1]].append(cs[0])

    def bfs_dist(srcs):
        d = np.full(n, -1)
        q = deque(srcs)
        for s in srcs:
            d[s] = 0
        while q:
            u = q.popleft()
            for w in adj[u]:
                if d[w] < 0:
                    d[w] = d[u] + 1
                    q.append(w)
        return d

    seeds = [0]
    for _ in range(1, n_parts):  # farthest-point seeding
        d = bfs_dist(seeds)
        seeds.append(int(np.argmax(np.where(d < 0, n, d))))
    rank = np.full(n, -1, dtype=np.int32)
    fronts = []
    for r, s in enumerate(seeds):
        rank[s] = r
        fronts.append(deque([s]))
    size = [1] * n_parts
    target = [(n + n_parts - 1 - r) // n_parts for r in range(n_parts)]
    left = n - n_parts
    while left > 0:
        progressed = False
        for r in range(n_parts):
            if size[r] >= target[r]:
                continue
            got = False
            while fronts[r] and not got:
                u = fronts[r][0]
                for w in adj[u]:
                    if rank[w] < 0:
                        rank[w] = r
                        fronts[r].append(w)
                        size[r] += 1
                        left -= 1
                        got = progressed = True
                        break
                if not got:
                    fronts[r].popleft()
        if not progressed:  # enclosed leftovers: give them to the smallest neighbouring region
            for c in np.nonzero(rank < 0)[0]:
                nb = [rank[w] for w in adj[c] if rank[w] >= 0]
                if nb:
                    r = min(nb, key=lambda q: size[q])
                    rank[c] = r
                    size[r] += 1
                    left -= 1
            if left > 0 and not any(rank[w] >= 0 for c in np.nonzero(rank < 0)[0] for w in adj[c]):
                rank[rank < 0] = int(np.argmin(size))
                left = 0
    # diffusion: hand boundary cells from larger to smaller neighbouring regions until the sizes differ by at most one
    for _ in range(4 * n):
        moved = False
        for c in range(n):
            a = rank[c]
            best = None
            for w in adj[c]:
                b = rank[w]
                if b != a and size[a] > size[b] + 1 and (best is None or size[b] < size[best]):
                    best = b
            if best is not None and sum(1 for w in adj[c] if rank[w] == a) >= 1:
                rank[c] = best
                size[a] -= 1
                size[best] += 1
                moved = True
        if not moved or max(size) - min(size) <= 1:
            break
    return rank
