"""bench.py -- P1 constant-stencil apply throughput (GDoF/s) on cube_6el level 8, with roofline, V(3,3) time,
end-to-end (host buffers through the C-ABI) and the reference's CPU kernels as baseline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--level L] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...      (one rank per GPU)

One "step" = one Operator::apply(src, dst, level, Inner) over the whole mesh.  See DESIGN.md (Measurement).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# dram__bytes_read.sum + dram__bytes_write.sum of one hb_shuffle_kernel<0,4,128> launch at level 8 on cube_6el
# (ncu --set full capture, profiles/r1_c_apply_final_ncu_full_summary.txt): 143.2 MB + 93.1 MB
NCU_TRAFFIC_INTERIOR_L8 = 236.3e6
BYTES_PER_DOF = {"apply": 16.0, "jacobi": 24.0, "sor": 24.0}  # algorithmic: src+dst / src+rhs+dst


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--level", type=int, default=8)
    p.add_argument("--impl", default="ours")
    p.add_argument("--no-cpu-baseline", action="store_true")
    return p.parse_args()


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """samples SM clock + throttle reasons through NVML during the timed region"""

    def __init__(self, index):
        self.samples, self.reasons, self.stop_flag, self.max_mhz = [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        nv = self.nv
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.02)

    def __enter__(self):
        if self.nv:
            self.t.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        if self.nv:
            self.t.join()

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own generated macro-cell apply kernel (oracle/_ref)
# ------------------------------------------------------------------------------------------------
def cpu_reference_apply(level, n_cells=6, warmup=1, steps=3, threads=None, budget_s=120.0):
    """Times apply_3D_macrocell_vertexdof_to_vertexdof_replace (compiled from the reference tree into
    oracle/_ref) on n_cells macro-cells of `level`, one thread per macro-cell (the reference parallelises
    over macro-primitives).  Only the macro-cell kernel is timed: the reference's macro-face/edge/vertex
    kernels and ghost-layer copies (~36 % of its apply) are left out, which favours the CPU.
    One step = one apply over all n_cells macro-cells; `warmup` untimed steps, then `steps` timed ones (mean)."""
    import ctypes as C
    so = os.path.join(ROOT, "oracle", "_ref", "libhyteg_ref_kernels.so")
    if not os.path.exists(so):
        return None
    dll = C.CDLL(so)
    dp = C.POINTER(C.c_double)
    dll.ref_apply_macrocell.argtypes = [dp, dp, C.c_int, dp, C.c_int]
    n = (2 ** level + 1) * (2 ** level + 2) * (2 ** level + 3) // 6
    N = 2 ** level
    interior = (N - 1) * (N - 2) * (N - 3) // 6
    threads = threads or min(n_cells, os.cpu_count() or 1)
    rng = np.random.default_rng(0)
    src = [rng.standard_normal(n) for _ in range(n_cells)]
    dst = [np.zeros(n) for _ in range(n_cells)]
    w = np.array([8.0] + [-0.5] * 14)

    def work(ids):
        for i in ids:
            dll.ref_apply_macrocell(dst[i].ctypes.data_as(dp), src[i].ctypes.data_as(dp), level, w.ctypes.data_as(dp), 0)

    def sweep():
        ts = [threading.Thread(target=work, args=(range(t, n_cells, threads),)) for t in range(threads)]
        [t.start() for t in ts]
        [t.join() for t in ts]

    t_start = time.perf_counter()
    for _ in range(warmup):
        sweep()
    times = []
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        sweep()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_start > budget_s:
            break
    mean = float(np.mean(times))
    return {"value": interior * n_cells / mean / 1e9, "unit": "GDoF/s", "cores": threads, "kind": "reference",
            "sample": f"reference macro-cell apply kernel only, {n_cells} macro-cells at level {level} (the whole workload), "
                      f"mean of {len(times)} applies after {warmup} warm-up",
            "seconds": mean, "steps": len(times)}


def run_reference(args, rank):
    """Reference arm: the reference's own generated apply kernel on the host cores, on this arm's workload (6 macro-cells per GPU
    of the job), args.warmup untimed and args.steps timed applies.  Rank 0 alone runs it."""
    if rank != 0:
        return
    base = cpu_reference_apply(args.level, n_cells=6 * max(1, args.gpus), warmup=args.warmup, steps=args.steps)
    if base is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"}))
        return
    v = base["value"]
    n_steps = base.pop("steps")
    workload = (f"cube_6el level {args.level} P1 Laplace constant-stencil apply" if args.gpus <= 1 else
                f"box of {args.gpus} cubes (6 macro-tets each) level {args.level} P1 Laplace constant-stencil apply") + " (CPU reference kernels)"
    line = {"impl": "reference", "metric": "P1 constant-stencil apply throughput", "value": v, "unit": "GDoF/s", "n_gpus": args.gpus,
            "steps": n_steps, "warmup": args.warmup, "ms_per_step": 1e3 * base.pop("seconds"), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload},
            "cpu_baseline": base, "e2e": {"value": v, "unit": "GDoF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import hyteg_b200 as hb
    from hyteg_b200 import api, meshes

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = hb.get_lib()
    lib.check(lib.dll.hb_device_init(local))
    stream = torch.cuda.current_stream()
    lib.check(lib.dll.hb_set_stream(stream.cuda_stream))

    L = args.level
    for kv in filter(None, os.environ.get("HB_TUNING", "").split(",")):  # A/B experiments: HB_TUNING="shuffle_depth=105,overlap=0"
        k, val = kv.split("=")
        lib.dll.hb_set_tuning(k.encode(), int(val))
    from hyteg_b200 import comm
    if world > 1:
        # weak scaling: a box of `world` unit cubes (6 macro-tets each), macro-cells partitioned in contiguous blocks so that
        # every rank owns one cube; the ghost layers on the faces between the cubes travel over NVLink inside the step: remote
        # stores into the neighbours' peer-memory windows (HB_P2P=0: pack -> ncclSend/ncclRecv -> unpack)
        v, c = meshes.cuboid(world, 1, 1, hi=(float(world), 1.0, 1.0))
        comm.init_nccl(lib)
    else:
        v, c = meshes.cube_6el()
    st = api.PrimitiveStorage(v, c, rank, world)
    st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    A = api.P1ConstantLaplaceOperator(st, 0, L)
    src, dst, rhs = (api.P1Function(n, st, 0, L) for n in ("src", "dst", "rhs"))
    n_arr = st.num_local_cells * st.cell_array_size(L)
    ndof = st.num_global_dofs(L)  # global: all ranks together
    host = torch.empty(n_arr, dtype=torch.float64).pin_memory()
    host.copy_(torch.from_numpy(np.random.default_rng(rank).standard_normal(n_arr)))
    hout = torch.empty(n_arr, dtype=torch.float64).pin_memory()
    hp = host.data_ptr()
    import ctypes as C
    dp = C.POINTER(C.c_double)
    # consistent copies of the interface DoFs: interpolate a smooth field, then perturb nothing (synthetic data)
    src.interpolate(lambda x, y, z: np.sin(3 * x) * np.cos(2 * y) + z * z, L, hb.All)
    rhs.interpolate(lambda x, y, z: np.cos(x + y) * z, L, hb.All)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps

    smooth = hb.SMOOTH_FLAG

    def apply_step():
        if world > 1:
            src.device_ptr(L)  # marks the mirrored ghost layers of src stale: every step pays its ghost exchange
        A.apply(src, dst, L, hb.Inner)

    W = max(args.warmup, 3)
    with ClockSampler(local) as clk:
        ms_apply = timed(apply_step, args.steps, W)
    ex0 = comm.exchange_stats(lib)
    # the dominant kernel alone (interior points, hb_shuffle_kernel): tuning knob apply_parts = 1 launches only that kernel
    lib.dll.hb_set_tuning(b"apply_parts", 1)
    ms_interior = timed(lambda: A.apply(src, dst, L, hb.Inner), args.steps, W)
    lib.dll.hb_set_tuning(b"apply_parts", 3)
    n_interior = st.num_local_cells * ((2 ** L - 1) * (2 ** L - 2) * (2 ** L - 3) // 6)
    ms_jac = timed(lambda: A.smooth_jac(dst, rhs, src, 0.66, L, smooth), args.steps, W)
    ms_sor = timed(lambda: A.smooth_sor(dst, rhs, 1.0, L, smooth), max(args.steps // 5, 2), 3)

    # V(3,3) cycle, Gauss-Seidel smoother, CG on level 0
    x, b = api.P1Function("x", st, 0, L), api.P1Function("b", st, 0, L)
    x.interpolate(1.0, L, hb.DirichletBoundary)
    gmg = api.GeometricMultigridSolver(st, api.GaussSeidelSmoother(st), api.CGSolver(st, 0, 0, 50, 1e-16, 1e-16), 0, L, 3, 3)
    ms_v = timed(lambda: gmg.solve(A, x, b, L), max(args.steps // 10, 2), 2)
    # same cycle with the fused weighted-Jacobi sweep as smoother (one kernel pair per sweep instead of ten SOR phases)
    gmg_j = api.GeometricMultigridSolver(st, api.WeightedJacobiSmoother(st, 0, L, 0.66), api.CGSolver(st, 0, 0, 50, 1e-16, 1e-16), 0, L, 3, 3)
    ms_vj = timed(lambda: gmg_j.solve(A, x, b, L), max(args.steps // 10, 2), 2)

    # end to end through the C-ABI with pinned HOST buffers: H2D src, apply, D2H dst every step.  n_lanes requests are kept in
    # flight on as many streams (own functions and host buffers each), so uploads overlap downloads (PCIe is full duplex);
    # every step still moves its full input and output.
    n_lanes = int(os.environ.get("HB_E2E_LANES", "2"))
    lanes = [(stream, src, dst, hout)]
    for k in range(1, n_lanes):
        lanes.append((torch.cuda.Stream(), api.P1Function("src%d" % k, st, L, L), api.P1Function("dst%d" % k, st, L, L),
                      torch.empty(n_arr, dtype=torch.float64).pin_memory()))
    n_e2e = max(args.steps // 5, 4) // n_lanes * n_lanes + n_lanes

    def e2e_run(steps):
        for k in range(steps):
            sk, fs, fd, ho = lanes[k % n_lanes]
            lib.check(lib.dll.hb_set_stream(sk.cuda_stream))
            lib.check(lib.dll.hb_p1function_upload_async(fs.h, L, C.cast(hp, dp)))
            A.apply(fs, fd, L, hb.Inner)
            lib.check(lib.dll.hb_p1function_download_async(fd.h, L, C.cast(ho.data_ptr(), dp)))
        lib.check(lib.dll.hb_set_stream(stream.cuda_stream))

    e2e_run(2 * n_lanes)
    barrier()
    t0 = time.perf_counter()
    e2e_run(n_e2e)
    torch.cuda.synchronize()
    ms_e2e = (time.perf_counter() - t0) * 1e3 / n_e2e
    if world > 1:
        t = torch.tensor([ms_e2e], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item())

    if rank == 0:
        peak, which = measured_peak()
        gd = lambda ms: ndof / (ms * 1e-3) / 1e9
        ndof_rank = ndof / world
        achieved = BYTES_PER_DOF["apply"] * n_interior / (ms_interior * 1e-3) / 1e9
        whole = BYTES_PER_DOF["apply"] * ndof_rank / (ms_apply * 1e-3) / 1e9
        npeers, nsend, nrecv = st.exchange_info(L)
        xmode = "NVLink peer-memory windows, one kernel per exchange" if (world > 1 and comm.p2p_active(lib)[0]) else "NCCL send/recv"
        line = {"metric": "P1 constant-stencil apply throughput", "value": gd(ms_apply), "unit": "GDoF/s", "n_gpus": world,
                "steps": args.steps, "warmup": W, "ms_per_step": ms_apply, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"cube_6el level {L} P1 Laplace constant-stencil apply, {int(ndof_rank)} DoFs per GPU" if world == 1 else
                           f"box of {world} cubes (6 macro-tets each) level {L}, partitioned one cube per GPU, {int(ndof_rank)} DoFs per GPU, "
                           f"ghost exchange of {8 * nsend} B with {npeers} neighbour rank(s) inside every step ({xmode})",
                           "l2": "src+dst arrays %.0f MB > 126 MB L2" % (2 * 8 * n_arr / 1e6),
                           "multi_gpu": f"macro-cells partitioned, ghost layers over {xmode}" if world > 1 else "single"},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": NCU_TRAFFIC_INTERIOR_L8 if L == 8 else None, "peak_source": which,
                             "kernel": "hb_shuffle_kernel<APPLY_REPLACE,4,128> (interior points of all local macro-cells, one launch)",
                             "algorithmic_bytes_per_launch": BYTES_PER_DOF["apply"] * n_interior, "us_per_launch": 1e3 * ms_interior,
                             "share_of_step": ms_interior / ms_apply,
                             "whole_apply": {"achieved": whole, "frac": whole / peak,
                                             "note": "16 B x all DoFs / apply time (interior + interface kernels, overlapped on two streams)"}},
                "also": {"jacobi_gdofs": gd(ms_jac), "jacobi_frac": BYTES_PER_DOF["jacobi"] * ndof_rank / (ms_jac * 1e-3) / 1e9 / peak,
                         "sor_gdofs": gd(ms_sor), "sor_frac": BYTES_PER_DOF["sor"] * ndof_rank / (ms_sor * 1e-3) / 1e9 / peak,
                         "vcycle_3_3_gs_ms": ms_v, "vcycle_3_3_jacobi_ms": ms_vj, "ghost_exchanges_in_apply_region": ex0[0]},
                "e2e": {"value": gd(ms_e2e), "unit": "GDoF/s", "h2d_bytes_per_step": 8 * n_arr, "d2h_bytes_per_step": 8 * n_arr,
                        "note": "%d requests in flight on %d streams (uploads overlap downloads, PCIe full duplex)" % (n_lanes, n_lanes)},
                "gpu_launches": 2 * args.steps + (2 * args.steps if world > 1 else 0), "clocks": clk.summary()}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_reference_apply(min(L, 8), warmup=1, steps=10)
            if line["cpu_baseline"]:
                line["cpu_baseline"].pop("steps", None)
        print(json.dumps(line))
    if world > 1:
        lib.check(lib.dll.hb_nccl_finalize())
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
