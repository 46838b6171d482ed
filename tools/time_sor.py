"""SOR sweep variants at one level (CUDA events): four colours vs line ordering, cell-wise launches, staging knobs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hyteg_b200 as hb
from hyteg_b200 import api, meshes
L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
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
INT_MIN = -2147483648
def run(label, **kv):
    for k, val in kv.items(): lib.dll.hb_set_tuning(k.encode(), val)
    st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    A = api.P1ConstantLaplaceOperator(st, L, L)
    x, b = api.P1Function("x", st, L, L), api.P1Function("b", st, L, L)
    x.interpolate(lambda X, Y, Z: X * Y + Z, L, hb.All)
    print("L%d %-40s fwd %.1f us  bwd %.1f us" % (L, label, t(lambda: A.smooth_sor(x, b, 1.0, L, hb.SMOOTH_FLAG)),
          t(lambda: A.smooth_sor(x, b, 1.0, L, hb.SMOOTH_FLAG, True))), flush=True)
    for k in kv: lib.dll.hb_set_tuning(k.encode(), INT_MIN)
    del A, x, b, st
run("default (four colours, coop six phases, direct)")
run("stage all", sor_stage_all=1)
run("coop small only (r1 phases)", sor_coop=2)
run("no coop", sor_coop=0)
run("depth 3", shuffle_depth=3)
run("depth 5", shuffle_depth=5)
