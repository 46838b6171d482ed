"""apply = interior kernel + interface kernel: each alone, serial and overlapped (CUDA events), on the cube and on the shell.
usage: time_apply_breakdown.py [LEVEL] [cube|shell]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hyteg_b200 as hb
from hyteg_b200 import api, meshes
L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
which = sys.argv[2] if len(sys.argv) > 2 else "cube"
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(0))
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
v, c = meshes.cube_6el() if which == "cube" else meshes.spherical_shell(2, 2, 0.5, 1.0)
INT_MIN = -2147483648
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(n): fn()
    b.record(s); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
A = api.P1ConstantLaplaceOperator(st, L, L)
src, dst = api.P1Function("s", st, L, L), api.P1Function("d", st, L, L)
src.interpolate(lambda X, Y, Z: X * Y + Z, L, hb.All)
nd = st.num_global_dofs(L)
def run(label, **kv):
    for k, val in kv.items(): lib.dll.hb_set_tuning(k.encode(), val)
    us = t(lambda: A.apply(src, dst, L, hb.All))
    print("%s L%d %-44s %8.1f us  %6.1f GDoF/s  %.3f of 6538.6 GB/s" % (which, L, label, us, nd / us * 1e-3, nd * 16 / us * 1e-3 / 6538.6), flush=True)
    for k in kv: lib.dll.hb_set_tuning(k.encode(), INT_MIN)
run("interior only", apply_parts=1)
run("interface only", apply_parts=2)
run("both, serial", overlap=0)
run("both, two streams (default)")
rhs = api.P1Function("r", st, L, L)
us = t(lambda: A.smooth_jac(dst, rhs, src, 0.66, L, hb.SMOOTH_FLAG))
print("%s L%d jacobi %8.1f us %6.1f GDoF/s" % (which, L, us, nd / us * 1e-3))
