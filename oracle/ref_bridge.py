"""TEST INFRASTRUCTURE ONLY: Python side of the reference driver (oracle/_ref/ref_dump, built by oracle/build_ref.py from
the reference's own sources).  Functions are handed over as point clouds (x, y, z, value) so that no assumption about the
reference's internal primitive numbering is needed; see oracle/ref_dump.cpp."""
import os
import subprocess
import tempfile

import numpy as np
from scipy.spatial import cKDTree

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DUMP = os.path.join(HERE, "_ref", "ref_dump")


def available():
    return os.path.exists(REF_DUMP)


def write_msh(v, c, path):
    """Gmsh 2.2 ASCII, tetrahedra only (MeshInfo::fromGmshFile reads this)"""
    with open(path, "w") as f:
        f.write("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n%d\n" % len(v))
        for i, p in enumerate(v):
            f.write("%d %.17g %.17g %.17g\n" % (i + 1, p[0], p[1], p[2]))
        f.write("$EndNodes\n$Elements\n%d\n" % len(c))
        for i, t in enumerate(c):
            f.write("%d 4 2 0 1 %d %d %d %d\n" % (i + 1, t[0] + 1, t[1] + 1, t[2] + 1, t[3] + 1))
        f.write("$EndElements\n")


def min_spacing(v, c, L):
    e = min(np.linalg.norm(v[t[a]] - v[t[b]]) for t in c for a in range(4) for b in range(a + 1, 4))
    return e / (1 << L)


class Session:
    """one macro mesh; clouds are exchanged through a temporary directory"""

    def __init__(self, v, c):
        self.v, self.c = np.asarray(v, dtype=np.float64), np.asarray(c)
        self.dir = tempfile.mkdtemp(prefix="refdump_")
        self.msh = os.path.join(self.dir, "mesh.msh")
        write_msh(self.v, self.c, self.msh)
        self.n = 0

    def cloud(self, xyz, val):
        self.n += 1
        p = os.path.join(self.dir, "c%d.bin" % self.n)
        np.concatenate([xyz, np.asarray(val, dtype=np.float64)[:, None]], axis=1).astype(np.float64).tofile(p)
        return p

    def run(self, cmd, level, **kv):
        """returns (results dict name -> list of floats, output cloud as (xyz, val) or None)"""
        out = os.path.join(self.dir, "out%d.bin" % self.n)
        if os.path.exists(out):
            os.remove(out)
        h = 0.4 * min_spacing(self.v, self.c, level)
        args = [REF_DUMP, cmd, "mesh=" + self.msh, "level=%d" % level, "h=%.17g" % h, "out=" + out]
        args += ["%s=%s" % (k, ("%.17g" % x) if isinstance(x, float) else x) for k, x in kv.items()]
        r = subprocess.run(args, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("ref_dump failed: %s\n%s\n%s" % (" ".join(args), r.stdout[-2000:], r.stderr[-2000:]))
        res = {}
        for line in r.stdout.splitlines():
            if line.startswith("RESULT "):
                _, k, x = line.split()
                res.setdefault(k, []).append(float(x))
        cl = None
        if os.path.exists(out):
            d = np.fromfile(out, dtype=np.float64).reshape(-1, 4)
            cl = (d[:, :3], d[:, 3])
        return res, cl


def to_global(xyz_global, cloud, tol):
    """cloud (xyz, val) with one record per macro-cell array entry -> vector over the unique DoFs at xyz_global;
    checks that all copies of a DoF carry the same value (the reference's ghost layers are consistent after communicate)"""
    xyz, val = cloud
    tree = cKDTree(xyz_global)
    d, idx = tree.query(xyz)
    assert d.max() < tol, d.max()
    out = np.full(len(xyz_global), np.nan)
    out[idx] = val
    assert not np.isnan(out).any()
    assert np.array_equal(out[idx], val), "copies of a DoF disagree in the reference dump"
    return out


def p2_apply(session, o, x, level, form="divkgrad", dst=None, flag=1, update=0, shell=0, **kw):
    """P2ElementwiseOperator< form >::apply of the reference on the P2 vector x (oracle numbering: vertex DoFs, then edge DoFs);
    inputs travel as ONE cloud in PHYSICAL coordinates (P2Function::interpolate evaluates at the blended position), the result comes
    back as a vertex cloud and an edge cloud in COMPUTATIONAL coordinates"""
    phys = o.dof_physical_coordinates()
    args = dict(src=session.cloud(phys, x), form=form, flag=flag, update=update, shell=shell, **kw)
    if dst is not None:
        args["dst"] = session.cloud(phys, dst)
    oute = os.path.join(session.dir, "oute%d.bin" % session.n)
    res, cl = session.run("p2apply", level, oute=oute, **args)
    de = np.fromfile(oute, dtype=np.float64).reshape(-1, 4)
    tol = 0.05 * min_spacing(session.v, session.c, level)
    v = to_global(o.xyz[: o.nv], cl, tol)
    e = to_global(o.xyz[o.nv:], (de[:, :3], de[:, 3]), tol)
    return np.concatenate([v, e])
