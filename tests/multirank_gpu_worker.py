"""Worker of tests/test_gpu_multirank.py: one rank (one GPU) of a world_size-2 NCCL job running the PRODUCT library.
Checks that the partitioned path (ghost exchange inside the library: one kernel over peer-memory windows, or with HB_P2P=0
pack kernel -> ncclSend/ncclRecv -> unpack kernel)
reproduces the one-rank run on the same mesh bit for bit, and the oracle within 1e-12."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import hyteg_b200 as hb  # noqa: E402
from hyteg_b200 import api, comm, meshes  # noqa: E402
from hyteg_b200.capi import All, Inner, DirichletBoundary  # noqa: E402
from oracle import p1_oracle as po  # noqa: E402
import parity_cases as pc  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    lib = hb.get_lib()
    lib.check(lib.dll.hb_device_init(local))
    comm.init_nccl(lib)
    for kv in filter(None, os.environ.get("HB_TUNING", "").split(",")):  # e.g. exchange_groups=0
        k, val = kv.split("=")
        lib.dll.hb_set_tuning(k.encode(), int(val))
    mesh = meshes.cuboid(world, 1, 1, hi=(float(world), 1.0, 1.0))
    own = lambda a: ~np.isnan(a)

    # oracle parity on a small level (generic kernels)
    L = 3
    s = pc.Setup(lib, mesh, L, L, rank, world)
    x, d0 = s.rand(L, 1), s.rand(L, 2)
    src, dst = s.fn("src"), s.fn("dst")
    s.put(src, x, L)
    s.put(dst, d0, L)
    s.A.apply(src, dst, L, Inner)
    got = s.get(dst, L)
    m = own(got)
    want = po.apply(s.o, x, d0, L, Inner)
    assert pc.rel(got[m], want[m]) < pc.TOL
    mi = s.o.mask(L, Inner)
    assert abs(src.dotGlobal(dst, L, Inner) - x[mi] @ want[mi]) < 1e-11 * abs(x[mi] @ want[mi])

    # level 5 (N = 32: shuffle interior kernel + interface kernel): partitioned run == one-rank run, bit for bit
    L = 5

    def run(r, w):
        t = api.PrimitiveStorage(mesh[0], mesh[1], r, w)
        t.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        A = api.P1ConstantLaplaceOperator(t, 0, L)
        fx, fb, fr = (api.P1Function(n, t, 0, L) for n in "xbr")
        fx.interpolate(lambda X, Y, Z: np.sin(3 * X) * np.cos(2 * Y) + Z * Z, L, All)
        fb.interpolate(lambda X, Y, Z: np.cos(X + Y) * Z, L, All)
        A.apply(fx, fr, L, Inner)
        ap = fr.download(L).copy()
        A.smooth_jac(fr, fb, fx, 0.66, L, pc.SMOOTH)
        A.smooth_sor(fr, fb, 1.0, L, pc.SMOOTH)
        A.smooth_sor(fr, fb, 1.0, L, pc.SMOOTH, backwards=True)
        so = fr.download(L).copy()
        fx.interpolate(0.0, L, Inner)
        fb.interpolate(0.0, L, All)
        gmg = api.GeometricMultigridSolver(t, api.GaussSeidelSmoother(t), api.CGSolver(t, 0, 0, 100, 1e-16, 1e-16), 0, L, 3, 3)
        res = []
        for _ in range(3):
            gmg.solve(A, fx, fb, L)
            A.residual(fx, fb, fr, L, Inner)
            res.append(np.sqrt(fr.dotGlobal(fr, L, Inner)))
        if w > 1:
            # the same cycles with levels 0 .. 3 on a replica of the whole mesh (gathered through the NCCL all-reduce) and level 0
            # solved directly: the same iteration up to rounding
            xs = fx.download(L).copy()
            fx.interpolate(0.0, L, Inner)
            rep = api.GeometricMultigridSolver(t, api.GaussSeidelSmoother(t), api.ReplicatedCoarseCycle(t, 3, "gs", 0.66, 3, 3, 0), 3, L, 3, 3)
            res_r = []
            for _ in range(3):
                rep.solve(A, fx, fb, L)
                A.residual(fx, fb, fr, L, Inner)
                res_r.append(np.sqrt(fr.dotGlobal(fr, L, Inner)))
            assert np.allclose(res_r, res, rtol=1e-7, atol=0), (res_r, res)
            assert np.abs(fx.download(L) - xs).max() <= 1e-9 * np.abs(xs).max()
        return t.local_cell_global_ids(), ap, so, fx.download(L).copy(), np.array(res)

    cells_p, ap_p, so_p, x_p, res_p = run(rank, world)
    calls, sent = comm.exchange_stats(lib)
    assert calls > 0 and sent > 0
    active, carried = comm.p2p_active(lib)
    if os.environ.get("HB_P2P", "1") != "0":  # one kernel per exchange over the neighbours' peer-memory windows
        assert active and carried == calls, (active, carried, calls)
    else:  # pack kernel -> ncclSend/ncclRecv -> unpack kernel
        assert not active and carried == 0
    cells_1, ap_1, so_1, x_1, res_1 = run(0, 1)
    n = len(ap_p) // len(cells_p)
    for i, c in enumerate(cells_p):
        j = cells_1.index(c)
        assert np.array_equal(ap_p[i * n:(i + 1) * n], ap_1[j * n:(j + 1) * n]), "apply differs from the one-rank run"
        assert np.array_equal(so_p[i * n:(i + 1) * n], so_1[j * n:(j + 1) * n]), "SOR differs from the one-rank run"
        assert np.abs(x_p[i * n:(i + 1) * n] - x_1[j * n:(j + 1) * n]).max() <= 1e-12 * np.abs(x_1).max()
    assert np.allclose(res_p, res_1, rtol=1e-9, atol=0), (res_p, res_1)
    assert res_p[-1] < 0.05 * res_p[0], res_p
    dist.barrier()
    print(f"OK rank {rank}/{world}: {len(cells_p)} cells, {calls} ghost exchanges, {8 * sent} bytes sent, residuals {res_p}")
    lib.check(lib.dll.hb_nccl_finalize())
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
