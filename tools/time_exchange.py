"""Ghost-exchange latency per level on N ranks (torchrun): pack kernel -> ncclSend/ncclRecv -> unpack kernel."""
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
comm_kw = {}
if len(sys.argv) > 1 and sys.argv[1] == "shell":
    v, c = meshes.spherical_shell(2, world + 1, 0.5, 1.0)
    st = api.PrimitiveStorage(v, c, rank, world, cell_rank=meshes.partition_radial_layers(v, c, world))
else:
    v, c = meshes.cuboid(world, 1, 1, hi=(float(world), 1.0, 1.0))
    st = api.PrimitiveStorage(v, c, rank, world)
st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
f = api.P1Function("f", st, 0, 8)
waves = [int(w) for w in os.environ.get("WAVES", "0").split(",")]
for L, w in [(L, w) for L in range(0, 9) for w in (waves if L >= 6 else waves[:1])]:
    lib.dll.hb_set_tuning(b"p2p_grid_waves", w)
    def step():
        f.device_ptr(L); f.communicate(L)
    for _ in range(5): step()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(50): step()
    b.record(s); torch.cuda.synchronize()
    if rank == 0:
        print(f"L{L}: grid waves {w or 8}: exchange {a.elapsed_time(b) / 50 * 1e3:7.1f} us, {8 * st.exchange_info(L)[1]} B sent")
lib.dll.hb_nccl_finalize(); dist.destroy_process_group()
