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



def ncu_traffic(level):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel from the committed ncu --set full capture
    (profiles/apply_traffic.json names the capture); None when no capture of this level is on file.  A number from a capture, not
    from this run: ncu replays one launch with a cold L2, so dirty lines still in L2 at the end make it read BELOW the algorithmic
    bytes."""
    try:
        with open(os.path.join(ROOT, "profiles", "apply_traffic.json")) as f:
            d = json.load(f)
        e = d.get(str(level))
        return (e["dram_read_bytes"] + e["dram_write_bytes"], e["source"]) if e else (None, None)
    except Exception:
        return None, None

BYTES_PER_DOF = {"apply": 16.0, "jacobi": 24.0, "sor": 24.0}  # algorithmic: src+dst / src+rhs+dst


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=50)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--level", type=int, default=8)
    p.add_argument("--impl", default="ours")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--mesh", default="auto", choices=["auto", "shell", "cubes"], help="auto: cube_6el on one GPU, the spherical shell on several")
    p.add_argument("--replicate-below", type=int, default=4,
                   help="more than one GPU: multigrid levels 0 .. this one run on a replica of the whole mesh on every rank (-1: all levels distributed)")
    p.add_argument("--p2-level", type=int, default=6, help="BASELINE config 3 (P2 div-k-grad with blending on the shell) level; 0: skip")
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
# reference arm / cpu baseline: the reference ITSELF (oracle/_ref/ref_dump = HyTeG built from /root/reference by oracle/build_ref.py)
# ------------------------------------------------------------------------------------------------
def workload_mesh(gpus, mesh="auto"):
    """macro mesh of the job: BASELINE configs[1] (cube_6el) on one GPU, configs[4] -- the icosahedral spherical shell with one radial
    layer of 60 macro-tets per GPU (meshSphericalShell(2, gpus + 1, 0.5, 1.0)) -- on several"""
    from hyteg_b200 import meshes
    if mesh == "cubes":
        return meshes.cuboid(gpus, 1, 1, hi=(float(gpus), 1.0, 1.0)), "cubes"
    if gpus <= 1 and mesh != "shell":
        return meshes.cube_6el(), "cube"
    return meshes.spherical_shell(2, gpus + 1, 0.5, 1.0), "shell"


def workload_name(level, gpus, mesh="auto"):
    """config.workload: the same string in both arms"""
    kind = "cubes" if mesh == "cubes" else ("cube" if gpus <= 1 and mesh != "shell" else "shell")
    if kind == "cube":
        return f"cube_6el level {level} P1 Laplace constant-stencil apply"
    if kind == "cubes":
        return f"box of {gpus} cubes (6 macro-tets each) level {level} P1 Laplace constant-stencil apply, one cube per GPU"
    return (f"spherical shell meshSphericalShell(2,{gpus + 1},0.5,1.0) ({60 * gpus} macro-tets) level {level} P1 Laplace constant-stencil apply, "
            f"one radial layer of 60 macro-tets per GPU")


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_reference_apply(level, gpus=1, budget_s=20.0, procs=None, mesh="auto"):
    """The reference's stock timed call -- P1ConstantLaplaceOperator::apply( src, dst, level, Inner )
    (apps/benchmarks/ApplyBenchmark/ApplyBenchmark.cpp:87; macro vertex/edge/face/cell kernels AND the six ghost-layer communicate
    phases in front of them) -- on the workload's macro mesh, one process per host core.  The reference parallelises with MPI ranks
    over macro-primitives; this build is serial (no MPI in the image), so every process runs the whole mesh on its own core and the
    throughputs are added: P independent ranks without inter-rank messages, which favours the CPU.  Each process times whole applies
    for budget_s seconds after one warm-up apply."""
    import subprocess, tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
    if not os.path.exists(exe):
        return None
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    from hyteg_b200 import meshes
    import ref_bridge
    # several GPUs: every process runs ONE radial layer of the shell (60 macro-tets, the share of one GPU) -- a bounded sample of the job
    (v, c), kind = workload_mesh(1 if mesh != "cubes" else gpus, "shell" if (gpus > 1 and mesh != "cubes") or mesh == "shell" else mesh)
    d = tempfile.mkdtemp(prefix="refbench_")
    msh = os.path.join(d, "mesh.msh")
    ref_bridge.write_msh(v, c, msh)
    procs = procs or host_cores()
    # memory: ~1 GB per process at level 8 per cube (3 functions with ghost layers + stencils)
    try:
        avail = os.sysconf("SC_AVPHYS_PAGES") * os.sysconf("SC_PAGE_SIZE")
        per = 0.2e9 * len(c) * 8 ** (level - 8)
        procs = max(1, min(procs, int(0.6 * avail / max(per, 1))))
    except Exception:
        pass
    cmd = [exe, "bench", "mesh=" + msh, "level=%d" % level, "minlevel=%d" % level, "what=apply", "iters=1000000", "seconds=%g" % budget_s]
    t0 = time.perf_counter()
    ps = [subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for _ in range(procs)]
    rates, calls, dofs = [], 0, 0.0
    for p in ps:
        out = p.communicate()[0]
        r = {l.split()[1]: float(l.split()[2]) for l in out.splitlines() if l.startswith("RESULT ")}
        if "seconds_per_call" in r:
            rates.append(r["dofs"] / r["seconds_per_call"])
            calls += int(r["calls"])
            dofs = r["dofs"]
    if not rates:
        return None
    return {"value": sum(rates) / 1e9, "unit": "GDoF/s", "cores": len(rates), "kind": "reference",
            "sample": f"P1ConstantLaplaceOperator::apply of the reference itself (all macro-primitive kernels + ghost-layer communication) on "
                      f"{workload_name(level, gpus, mesh)}" + (" -- every process runs the share of one GPU (one radial layer, 60 macro-tets)" if kind == "shell" and gpus > 1 else "") +
                      f": {len(rates)} serial processes, one per host core, each timing whole applies for "
                      f"{budget_s:g} s after one warm-up apply; throughputs added",
            "seconds": dofs / sum(rates), "steps": calls, "wall_s": time.perf_counter() - t0,
            "per_core_gdofs": float(np.mean(rates)) / 1e9}


def run_reference(args, rank):
    """Reference arm: the reference's own apply on the host cores, on this arm's workload.  Rank 0 alone runs it."""
    if rank != 0:
        return
    base = cpu_reference_apply(args.level, gpus=max(1, args.gpus), budget_s=max(5.0, min(60.0, 0.5 * args.steps)), mesh=args.mesh)
    if base is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_dump not built (needs /root/reference at build time)"}))
        return
    v = base["value"]
    n_steps = base.pop("steps")
    line = {"impl": "reference", "metric": "P1 constant-stencil apply throughput", "value": v, "unit": "GDoF/s", "n_gpus": args.gpus,
            "steps": n_steps, "warmup": 1, "ms_per_step": 1e3 * base.pop("seconds"), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.level, args.gpus, args.mesh)},
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
    try:  # pinned host buffers are first-touched on the NUMA node of this rank's GPU
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1} & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass
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
    # weak scaling (BASELINE configs[4]): the spherical shell with one radial layer of 60 macro-tets per GPU (a rank's neighbours are the
    # layers below and above it: two whole spherical surfaces of ghost data); the ghost layers travel over NVLink inside the step:
    # remote stores into the neighbours' peer-memory windows (HB_P2P=0: pack -> ncclSend/ncclRecv -> unpack)
    (v, c), mesh_kind = workload_mesh(world, args.mesh)
    cell_rank = None
    if world > 1:
        comm.init_nccl(lib, segment_bytes=256 << 20)  # the shell's interfaces are whole spherical surfaces: several million doubles per pair at level 8
        if mesh_kind == "shell":
            cell_rank = meshes.partition_radial_layers(v, c, world)
    st = api.PrimitiveStorage(v, c, rank, world, cell_rank=cell_rank)
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
    # more than one rank: the levels 0 .. LREP of the cycle run on a replica of the whole mesh, redundantly on every rank -- the same
    # operations without their ghost exchanges (~ 40 per level and cycle, three all-reduces per CG iteration); knob --replicate-below
    LREP = min(args.replicate_below, L - 1) if world > 1 else -1
    if LREP >= 0:
        # level 0 (12 (gpus + 1) macro vertices) by dense Cholesky on the host instead of <= 50 CG iterations with three host-read dot
        # products each (cg_max_iter = 0; the reference's counterpart as coarse-grid solver is PETScLUSolver)
        coarse_gs = api.ReplicatedCoarseCycle(st, LREP, "gs", 0.66, 3, 3, 0)
        coarse_j = api.ReplicatedCoarseCycle(st, LREP, "jacobi", 0.66, 3, 3, 0)
    else:
        coarse_gs = coarse_j = None
    gmg = api.GeometricMultigridSolver(st, api.GaussSeidelSmoother(st), coarse_gs or api.CGSolver(st, 0, 0, 50, 1e-16, 1e-16), max(LREP, 0), L, 3, 3)
    ms_v = timed(lambda: gmg.solve(A, x, b, L), max(args.steps // 10, 2), 2)
    # same cycle with the fused weighted-Jacobi sweep as smoother (one kernel pair per sweep instead of ten SOR phases)
    gmg_j = api.GeometricMultigridSolver(st, api.WeightedJacobiSmoother(st, 0, L, 0.66), coarse_j or api.CGSolver(st, 0, 0, 50, 1e-16, 1e-16), max(LREP, 0), L, 3, 3)
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

    # end-to-end SOLVE, the unit this path's callers see: upload b, V(3,3) cycles until the residual has dropped by 1e-8, download x
    hb_host = torch.empty(n_arr, dtype=torch.float64).pin_memory()
    rhs.interpolate(lambda X, Y, Z: np.sin(2 * X) * np.cos(3 * Y) * (1 + Z), L, hb.All)
    hb_host.copy_(torch.from_numpy(rhs.download(L)))
    r = api.P1Function("r", st, L, L)

    def e2e_solve():
        lib.check(lib.dll.hb_p1function_upload_async(b.h, L, C.cast(hb_host.data_ptr(), dp)))
        x.interpolate(0.0, L, hb.All)
        A.residual(x, b, r, L, hb.Inner)
        r0 = np.sqrt(r.dotGlobal(r, L, hb.Inner))
        cycles, res = 0, r0
        while res > 1e-8 * r0 and cycles < 30:
            gmg.solve(A, x, b, L)
            A.residual(x, b, r, L, hb.Inner)
            res = np.sqrt(r.dotGlobal(r, L, hb.Inner))
            cycles += 1
        lib.check(lib.dll.hb_p1function_download_async(x.h, L, C.cast(hout.data_ptr(), dp)))
        torch.cuda.synchronize()
        return cycles, res / r0

    e2e_solve()
    barrier()
    t0 = time.perf_counter()
    solve_cycles, solve_red = e2e_solve()
    ms_solve = (time.perf_counter() - t0) * 1e3
    if world > 1:
        t = torch.tensor([ms_solve], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_solve = float(t.item())

    # more than one rank: the partitioned apply and one V-cycle must equal the one-rank run bit for bit (small level, every rank
    # checks its own macro-cells against an unpartitioned storage on its own GPU) -- the driver's 1-GPU box skips the 2-GPU tests
    parity_ok = None
    if world > 1:
        Lc = 4
        st1 = api.PrimitiveStorage(v, c, 0, 1)  # the whole macro mesh on this GPU
        st1.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        ok = True
        A1, Ap = api.P1ConstantLaplaceOperator(st1, 0, Lc), api.P1ConstantLaplaceOperator(st, 0, Lc)
        f1 = [api.P1Function(n, st1, 0, Lc) for n in "xyb"]
        fp = [api.P1Function(n, st, 0, Lc) for n in "xyb"]
        expr = lambda X, Y, Z: np.sin(5 * X) * np.cos(3 * Y) + Z * Z
        for fs, stx, Ax in ((f1, st1, A1), (fp, st, Ap)):
            fs[0].interpolate(expr, Lc, hb.All)
            Ax.apply(fs[0], fs[1], Lc, hb.Inner)
            g = api.GeometricMultigridSolver(stx, api.GaussSeidelSmoother(stx), api.CGSolver(stx, 0, 0, 50, 1e-16, 1e-16), 0, Lc, 3, 3)
            fs[0].interpolate(1.0, Lc, hb.DirichletBoundary)
            g.solve(Ax, fs[0], fs[2], Lc)
        csz = st.cell_array_size(Lc)
        mine = st.local_cell_global_ids()
        for k in (0, 1):
            full, part = f1[k].download(Lc), fp[k].download(Lc)
            for i, gcell in enumerate(mine):
                a_, b_ = full[gcell * csz:(gcell + 1) * csz], part[i * csz:(i + 1) * csz]
                # k = 1: the apply result, bit for bit; k = 0: the iterate after one V-cycle -- its coarse-grid CG sums its dot
                # products rank by rank, so the last bits may differ
                ok = ok and (np.array_equal(a_, b_) if k == 1 else bool(np.abs(a_ - b_).max() <= 1e-12 * max(np.abs(a_).max(), 1e-300)))
        t = torch.tensor([1.0 if ok else 0.0], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        parity_ok = bool(t.item() > 0.5)

    # BASELINE config 3: P2 elementwise div-k-grad with IcosahedralShellMap blending on the thick spherical shell (240 macro-tets)
    p2 = None
    if world == 1 and args.p2_level > 0:
        P2L = args.p2_level
        t0 = time.perf_counter()
        vs, cs = meshes.spherical_shell(3, 2, 0.5, 1.0)
        sts = api.PrimitiveStorage(vs, cs)
        sts.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        api.setIcosahedralShellMap(sts)
        ps, pd, pk = (api.P2Function(n, sts, P2L, P2L) for n in ("src", "dst", "k"))
        ps.interpolate(1.0, P2L, hb.All)
        rng = np.random.default_rng(1)
        pk.interpolate(1.5, P2L, hb.All)
        Ap2 = api.P2ElementwiseBlendingDivKGradOperator(sts, P2L, P2L, pk)
        Ap2.apply(ps, pd, P2L, hb.Inner)
        torch.cuda.synchronize()
        setup_s = time.perf_counter() - t0
        ms_p2 = timed(lambda: Ap2.apply(ps, pd, P2L, hb.Inner), max(args.steps // 10, 3), 2)
        n_tet = len(cs) * 8 ** P2L
        flop = n_tet * 11 * 330.0  # 11 quadrature points x ~330 FP64 flop (geometry + the two contractions), DESIGN.md section 9
        p2 = {"workload": f"P2 elementwise div-k-grad, IcosahedralShellMap blending, meshSphericalShell(3,2,0.5,1.0) level {P2L}, k as a P2 function",
              "ms_per_apply": ms_p2, "micro_tets": n_tet, "Gtets_per_s": n_tet / ms_p2 / 1e6, "fp64_tflops_estimate": flop / ms_p2 / 1e9,
              "setup_s": setup_s}
        del Ap2, ps, pd, pk, sts

    npeers, nsend, nrecv = st.exchange_info(L)
    npeers_max = npeers
    if world > 1:
        t = torch.tensor([float(npeers), float(nsend)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        npeers_max, nsend = int(t[0].item()), int(t[1].item())
    if rank == 0:
        peak, which = measured_peak()
        gd = lambda ms: ndof / (ms * 1e-3) / 1e9
        ndof_rank = ndof / world
        achieved = BYTES_PER_DOF["apply"] * n_interior / (ms_interior * 1e-3) / 1e9
        whole = BYTES_PER_DOF["apply"] * ndof_rank / (ms_apply * 1e-3) / 1e9
        xmode = "NVLink peer-memory windows, one kernel per exchange" if (world > 1 and comm.p2p_active(lib)[0]) else "NCCL send/recv"
        line = {"metric": "P1 constant-stencil apply throughput", "value": gd(ms_apply), "unit": "GDoF/s", "n_gpus": world,
                "steps": args.steps, "warmup": W, "ms_per_step": ms_apply, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(L, world, args.mesh), "dofs_per_gpu": int(ndof_rank), "dofs_total": int(ndof),
                           "exchange": None if world == 1 else f"ghost exchange of up to {8 * nsend} B per rank with up to {npeers_max} neighbour ranks inside every step ({xmode})",
                           "l2": "src+dst arrays %.0f MB > 126 MB L2" % (2 * 8 * n_arr / 1e6),
                           "multi_gpu": f"macro-cells partitioned, ghost layers over {xmode}" if world > 1 else "single"},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": ncu_traffic(L)[0], "traffic_source": ncu_traffic(L)[1], "peak_source": which,
                             "kernel": "hb_shuffle_kernel<APPLY_REPLACE,4,128> (interior points of all local macro-cells, one launch)",
                             "algorithmic_bytes_per_launch": BYTES_PER_DOF["apply"] * n_interior, "us_per_launch": 1e3 * ms_interior,
                             "share_of_step": ms_interior / ms_apply,
                             "whole_apply": {"achieved": whole, "frac": whole / peak,
                                             "note": "16 B x all DoFs / apply time (interior + interface kernels, overlapped on two streams)"}},
                "also": {"jacobi_gdofs": gd(ms_jac), "jacobi_frac": BYTES_PER_DOF["jacobi"] * ndof_rank / (ms_jac * 1e-3) / 1e9 / peak,
                         "sor_gdofs": gd(ms_sor), "sor_frac": BYTES_PER_DOF["sor"] * ndof_rank / (ms_sor * 1e-3) / 1e9 / peak,
                         "vcycle_3_3_gs_ms": ms_v, "vcycle_3_3_jacobi_ms": ms_vj, "vcycle_levels_replicated_on_every_rank": (list(range(0, LREP + 1)) if LREP >= 0 else []),
                         "vcycle_level0_solver": ("dense Cholesky on the replica" if LREP >= 0 else "CG, 50 iterations max, tolerance 1e-16"), "ghost_exchanges_in_apply_region": ex0[0],
                         "e2e_solve": {"ms": ms_solve, "vcycles": solve_cycles, "residual_reduction": solve_red, "gdofs_solved_per_s": ndof / ms_solve / 1e6,
                                       "what": "upload b (pinned host), V(3,3) Gauss-Seidel cycles until |r| <= 1e-8 |r0| (one host-read dot per cycle), download x"},
                         "gs_factor_gap": "multicolour vs the reference's lexicographic hybrid, V(3,3) factor over cycles 6..10: -3.5 % cube_6el L5, "
                                          "+2.6 % octahedron L4, -1.1 % three_tets L4 (tests/test_gpu_reference.py, DESIGN.md section 3)",
                         "parity_ok": parity_ok, "p2_config3": p2},
                "e2e": {"value": gd(ms_e2e), "unit": "GDoF/s", "h2d_bytes_per_step": 8 * n_arr, "d2h_bytes_per_step": 8 * n_arr,
                        "note": "%d requests in flight on %d streams (uploads overlap downloads, PCIe full duplex)" % (n_lanes, n_lanes)},
                "gpu_launches": 2 * args.steps + (2 * args.steps if world > 1 else 0), "clocks": clk.summary()}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_reference_apply(min(L, 8), budget_s=10.0, mesh=args.mesh)
            if line["cpu_baseline"]:
                line["cpu_baseline"].pop("steps", None)
        print(json.dumps(line))
    if world > 1:
        lib.check(lib.dll.hb_nccl_finalize())
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
