"""Small driver for ncu: a few apply / Jacobi / SOR launches on cube_6el at --level (default 8)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hyteg_b200 as hb
from hyteg_b200 import api, meshes

L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
lib = hb.get_lib()
lib.check(lib.dll.hb_device_init(0))
v, c = meshes.cube_6el()
st = api.PrimitiveStorage(v, c)
st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
A = api.P1ConstantLaplaceOperator(st, L, L)
src, dst, rhs = (api.P1Function(n, st, L, L) for n in ("src", "dst", "rhs"))
src.interpolate(lambda x, y, z: np.sin(x) + y * z, L, hb.All)
for _ in range(reps):
    A.apply(src, dst, L, hb.Inner)
for _ in range(reps):
    A.smooth_jac(dst, rhs, src, 0.66, L, hb.SMOOTH_FLAG)
for _ in range(reps):
    A.smooth_sor(src, rhs, 1.0, L, hb.SMOOTH_FLAG)
lib.check(lib.dll.hb_synchronize())
print("done", src.dotGlobal(src, L, hb.All))
