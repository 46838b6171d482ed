"""Per-level times of the building blocks of the V-cycle (CUDA events): SOR sweep, residual, restrict, prolongate."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hyteg_b200 as hb
from hyteg_b200 import api, comm, meshes
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(local))
if world > 1:  # torchrun: box of `world` cubes, one per rank (the bench's weak-scaling mesh)
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    comm.init_nccl(lib, segment_bytes=256 << 20)
for kv in filter(None, os.environ.get("HB_TUNING", "").split(",")):
    k, val = kv.split("="); lib.dll.hb_set_tuning(k.encode(), int(val))
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
LMAX = 8
if world == 1 and os.environ.get("MESH", "cube") == "cube":
    v, c = meshes.cube_6el()
    st = api.PrimitiveStorage(v, c, rank, world)
else:  # the bench's weak-scaling mesh: one radial layer of the shell per rank
    v, c = meshes.spherical_shell(2, world + 1, 0.5, 1.0)
    st = api.PrimitiveStorage(v, c, rank, world, cell_rank=meshes.partition_radial_layers(v, c, world))
st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
A = api.P1ConstantLaplaceOperator(st, 0, LMAX)
x, b, r = (api.P1Function(n, st, 0, LMAX) for n in "xbr")
R, P = api.P1toP1LinearRestriction(), api.P1toP1LinearProlongation()
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(n): fn()
    e.record(s); torch.cuda.synchronize()
    return a.elapsed_time(e) / n * 1e3
tot = 0.0
for L in range(1, LMAX + 1):
    sor = t(lambda: A.smooth_sor(x, b, 1.0, L, hb.SMOOTH_FLAG))
    res = t(lambda: A.residual(x, b, r, L, hb.Inner))
    rs = t(lambda: R.restrict(r, L, hb.SMOOTH_FLAG))
    pr = t(lambda: P.prolongateAndAdd(x, L - 1, hb.SMOOTH_FLAG))
    lvl = 6 * sor + res + rs + pr
    tot += lvl
    if rank == 0: print(f"L{L}: sor {sor:8.1f} us  residual {res:7.1f}  restrict {rs:7.1f}  prolongateAndAdd {pr:7.1f}  | V(3,3) share {lvl:8.1f} us")
if rank == 0: print(f"sum over levels {tot/1e3:.2f} ms (plus CG on level 0)")
# the coarse-grid solver and the whole cycle as the bench runs it
cg = api.CGSolver(st, 0, 0, 50, 1e-16, 1e-16)  # as in bench.py
b.interpolate(lambda X, Y, Z: X + Y * Z, 0, hb.All)
if rank == 0: print("level 0:", st.num_global_dofs(0), "DoFs")
def coarse():
    x.interpolate(0.0, 0, hb.Inner)
    cg.solve(A, x, b, 0)
import time
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20): coarse()
torch.cuda.synchronize()
if rank == 0: print(f"CG on level 0: {(time.perf_counter() - t0) / 20 * 1e6:.1f} us wall per solve, {cg.iterations()} iterations")
gmg = api.GeometricMultigridSolver(st, api.GaussSeidelSmoother(st), cg, 0, LMAX, 3, 3)
b.interpolate(lambda X, Y, Z: X + Y * Z, LMAX, hb.All)
for _ in range(2): gmg.solve(A, x, b, LMAX)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): gmg.solve(A, x, b, LMAX)
torch.cuda.synchronize()
if rank == 0: print(f"V(3,3) GS: {(time.perf_counter() - t0) / 10 * 1e3:.3f} ms wall per cycle")
