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
