"""ncu driver: interior-only apply at level 8 with tuning knobs given as key=value arguments."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hyteg_b200 as hb
from hyteg_b200 import api, meshes
L = 8
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(0))
for kv in sys.argv[1:]:
    k, v = kv.split("=")
    if k == "level":
        L = int(v)
    else:
        lib.dll.hb_set_tuning(k.encode(), int(v))
v, c = meshes.cube_6el()
st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
A = api.P1ConstantLaplaceOperator(st, L, L)
src, dst, rhs = api.P1Function("s", st, L, L), api.P1Function("d", st, L, L), api.P1Function("r", st, L, L)
src.interpolate(1.0, L, hb.All)
for _ in range(3):
    A.apply(src, dst, L, hb.All)
A.smooth_jac(dst, rhs, src, 0.66, L, hb.SMOOTH_FLAG)
lib.check(lib.dll.hb_synchronize())
print("done")
