"""Wall time of the building blocks of the coarse-grid CG on N ranks (torchrun), shell with one radial layer per rank."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import hyteg_b200 as hb
from hyteg_b200 import api, comm, meshes
local = int(os.environ.get("LOCAL_RANK", 0)); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(local)); comm.init_nccl(lib, segment_bytes=64 << 20)
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
v, c = meshes.spherical_shell(2, world + 1, 0.5, 1.0)
st = api.PrimitiveStorage(v, c, rank, world, cell_rank=meshes.partition_radial_layers(v, c, world)); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
for L in (0, 2, 4):
    A = api.P1ConstantLaplaceOperator(st, L, L)
    x, y, z = (api.P1Function(n, st, L, L) for n in "xyz")
    x.interpolate(lambda X, Y, Z: X * Y + Z, L, hb.All)
    def wall(fn, n=50):
        for _ in range(5): fn()
        torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
        for _ in range(n): fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n * 1e6
    r = {"dot": wall(lambda: x.dotGlobal(x, L, hb.Inner)),
         "apply (exchange)": wall(lambda: (x.device_ptr(L), A.apply(x, y, L, hb.Inner))),
         "apply (no exchange)": wall(lambda: A.apply(x, y, L, hb.Inner)),
         "exchange": wall(lambda: (x.device_ptr(L), x.communicate(L))),
         "add": wall(lambda: z.add([0.5], [x], L, hb.Inner)),
         "interpolate const": wall(lambda: z.interpolate(0.0, L, hb.All)),
         "sor": wall(lambda: A.smooth_sor(x, y, 1.0, L, hb.SMOOTH_FLAG))}
    def cg_like():  # one CG iteration's worth of operations, in CG's order
        A.apply(x, y, L, hb.Inner); x.dotGlobal(y, L, hb.Inner); z.add([0.5], [x], L, hb.Inner); z.add([-0.5], [y], L, hb.Inner)
        z.dotGlobal(z, L, hb.Inner); y.interpolate(0.0, L, hb.All); y.assign([1.0], [z], L, hb.Inner); z.dotGlobal(y, L, hb.Inner)
        x.assign([1.0, 0.3], [y, x], L, hb.Inner)
    r["one CG iteration's operations"] = wall(cg_like, 20)
    cg = api.CGSolver(st, L, L, 10, 1e-30, 1e-300)
    y.interpolate(lambda X, Y, Z: X + Y * Z, L, hb.All)
    def solve():
        z.interpolate(0.0, L, hb.Inner); cg.solve(A, z, y, L)
    r["CGSolver, 10 iterations"] = wall(solve, 5)
    if rank == 0: print(f"L{L} on {world} ranks (us wall):", ", ".join(f"{k} {v:.1f}" for k, v in r.items()), flush=True)
lib.dll.hb_nccl_finalize(); dist.destroy_process_group()
