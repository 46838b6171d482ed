"""apply / Jacobi with a ghost exchange in every step on N ranks (torchrun), shell with one radial layer per rank, level 8: how the grid of
the stand-alone exchange kernel (tuning p2p_grid_waves, in units of 148 CTAs) shares the SMs with the interior kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import hyteg_b200 as hb
from hyteg_b200 import api, comm, meshes
local = int(os.environ.get("LOCAL_RANK", 0)); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(local)); comm.init_nccl(lib, segment_bytes=256 << 20)
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
v, c = meshes.spherical_shell(2, world + 1, 0.5, 1.0)
st = api.PrimitiveStorage(v, c, rank, world, cell_rank=meshes.partition_radial_layers(v, c, world)); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
A = api.P1ConstantLaplaceOperator(st, L, L)
src, dst, rhs = (api.P1Function(n, st, L, L) for n in "sdr")
src.interpolate(lambda X, Y, Z: X * Y + Z, L, hb.All)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(n): fn()
    b.record(s); torch.cuda.synchronize()
    x = torch.tensor([a.elapsed_time(b) / n * 1e3], device="cuda"); dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return float(x)
def apply_stale():
    src.device_ptr(L); A.apply(src, dst, L, hb.All)
def jac_stale():
    src.device_ptr(L); A.smooth_jac(dst, rhs, src, 0.66, L, hb.SMOOTH_FLAG)
us0 = t(lambda: A.apply(src, dst, L, hb.All))
if rank == 0: print(f"L{L} {world} ranks: apply without exchange {us0:.1f} us")
for w in [int(x) for x in os.environ.get("WAVES", "1,2,3,4,8").split(",")]:
    lib.dll.hb_set_tuning(b"p2p_grid_waves", w)
    ua, uj = t(apply_stale), t(jac_stale)
    if rank == 0: print(f"   exchange kernel on {w} x 148 CTAs: apply {ua:.1f} us   jacobi {uj:.1f} us")
lib.dll.hb_nccl_finalize(); dist.destroy_process_group()
