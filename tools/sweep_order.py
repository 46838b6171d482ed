"""Interior-kernel sweep over item height / CTA pacing (run plain for times, or under ncu for DRAM bytes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hyteg_b200 as hb
from hyteg_b200 import api, meshes, capi
L = 8
if os.environ.get("HB_LIB"):  # experiment variant built by tools/build_variant.sh
    capi._default = capi.Lib(os.environ["HB_LIB"])
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(0))
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
v, c = meshes.cube_6el()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
def t(fn, n=reps):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(n): fn()
    b.record(s); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
lib.dll.hb_set_tuning(b"apply_parts", 1)
for ys in (32,):
    for sync in (0, 100, 200, 1000, 1100, 1200):  # here: order * 100 + cell_major * 1000 (cell_major inverted: 1000 = item-major)
        order, itemmajor = (sync // 100) % 10, sync // 1000
        lib.dll.hb_set_tuning(b"shuffle_ys", ys); lib.dll.hb_set_tuning(b"shuffle_sync", 0)
        lib.dll.hb_set_tuning(b"shuffle_order", order); lib.dll.hb_set_tuning(b"shuffle_cell_major", 1 - itemmajor)
        st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        A = api.P1ConstantLaplaceOperator(st, L, L)
        src, dst = api.P1Function("s", st, L, L), api.P1Function("d", st, L, L)
        src.interpolate(1.0, L, hb.All)
        r = []
        for depth in (4,):
            lib.dll.hb_set_tuning(b"shuffle_depth", depth)
            r.append("d%d %.1f us" % (depth, t(lambda: A.apply(src, dst, L, hb.All))))
        print("ys", ys, "sync", sync, *r, flush=True)
        del A, src, dst, st
