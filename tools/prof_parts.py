"""ncu driver: apply with only the interface part (apply_parts=2) or only the interior (1) at --level."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hyteg_b200 as hb
from hyteg_b200 import api, meshes
L = int(sys.argv[1]); parts = int(sys.argv[2])
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(0))
v, c = meshes.cube_6el()
st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
A = api.P1ConstantLaplaceOperator(st, L, L)
src, dst = api.P1Function("s", st, L, L), api.P1Function("d", st, L, L)
src.interpolate(lambda x, y, z: np.sin(x) + y * z, L, hb.All)
lib.dll.hb_set_tuning(b"apply_parts", parts)
for _ in range(3):
    A.apply(src, dst, L, hb.All)
lib.check(lib.dll.hb_synchronize())
print("done")
