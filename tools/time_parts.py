"""Times apply with interior only / interface only / both (tuning knob apply_parts) using CUDA events via torch.
usage: time_parts.py LEVEL [ys,ys,...] [order,...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hyteg_b200 as hb
from hyteg_b200 import api, meshes
L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
yss = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0]
orders = [int(v) for v in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0]
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(0))
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
v, c = meshes.cube_6el()
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(n): fn()
    b.record(s); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for ys in yss:
    for order in orders:
        lib.dll.hb_set_tuning(b"shuffle_ys", ys); lib.dll.hb_set_tuning(b"shuffle_order", order)
        st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        A = api.P1ConstantLaplaceOperator(st, L, L)
        src, dst, rhs = api.P1Function("s", st, L, L), api.P1Function("d", st, L, L), api.P1Function("r", st, L, L)
        src.interpolate(1.0, L, hb.All)
        for depth in (3, 4, 5, 7):
            lib.dll.hb_set_tuning(b"shuffle_depth", depth); lib.dll.hb_set_tuning(b"apply_parts", 1)
            print("ys", ys, "order", order, "depth", depth, "interior %.1f us" % t(lambda: A.apply(src, dst, L, hb.All)), flush=True)
        lib.dll.hb_set_tuning(b"shuffle_depth", 0)
        for parts in (2, 3):
            lib.dll.hb_set_tuning(b"apply_parts", parts)
            print("   apply_parts", parts, "%.1f us" % t(lambda: A.apply(src, dst, L, hb.All)))
        for ov, dep in ((1, 0), (1, 105), (1, 203)):
            lib.dll.hb_set_tuning(b"overlap", ov); lib.dll.hb_set_tuning(b"shuffle_depth", dep)
            print("   overlap", ov, "depth", dep, "apply %.1f us" % t(lambda: A.apply(src, dst, L, hb.All)))
        lib.dll.hb_set_tuning(b"overlap", 0); lib.dll.hb_set_tuning(b"shuffle_depth", 0)
        for depth in (3, 4, 5):
            lib.dll.hb_set_tuning(b"shuffle_depth", depth)
            print("   depth", depth, "jacobi %.1f us" % t(lambda: A.smooth_jac(dst, rhs, src, 0.66, L, hb.SMOOTH_FLAG)),
                  "sor %.1f us" % t(lambda: A.smooth_sor(dst, rhs, 1.0, L, hb.SMOOTH_FLAG), 5))
        lib.dll.hb_set_tuning(b"shuffle_depth", 0)
        del A, src, dst, rhs, st
