"""Worker of tests/test_multirank_gloo.py: one rank of a world_size-N gloo job on CPU.

Binds the TEST-ONLY emulation build, partitions the macro mesh over the ranks and checks that the partitioned
path (ghost exchange through the exchange callback) reproduces the oracle and, bit for bit, the one-rank run."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch.distributed as dist  # noqa: E402

from hyteg_b200 import api, capi, comm, meshes  # noqa: E402
from hyteg_b200.capi import All, Inner, DirichletBoundary  # noqa: E402
from oracle import p1_oracle as po  # noqa: E402
import parity_cases as pc  # noqa: E402


def own(a):
    return ~np.isnan(a)


def main():
    mesh_name = sys.argv[1]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    lib = capi.Lib(os.path.join(ROOT, "tests", "emul", "libhyteg_b200_emul.so"))
    comm.init_callbacks(lib)
    if mesh_name == "cuboid":
        mesh = meshes.cuboid(2, 1, 1)
    elif mesh_name == "shell":  # BASELINE config[4] family: icosahedral spherical shell (60 macro-cells), many neighbours per rank
        mesh = meshes.spherical_shell(2, 2)
    elif mesh_name == "shell_greedy":  # two radial layers, connected blocks from the greedy balancer: every rank has several neighbours
        mesh = meshes.spherical_shell(2, 3)
    else:
        mesh = meshes.golden(mesh_name)
    cell_rank = meshes.partition_greedy(mesh[1], world) if mesh_name == "shell_greedy" else None
    L = 3

    # ---- apply / Jacobi / BLAS-1 / transfer against the oracle ---------------------------------------------------
    s = pc.Setup(lib, mesh, L - 1, L, rank, world, cell_rank)
    assert 0 < s.st.num_local_cells < len(mesh[1])
    npeers, nsend, nrecv = s.st.exchange_info(L)
    assert 1 <= npeers <= world - 1 and nsend > 0 and nrecv > 0, (npeers, nsend, nrecv)
    x, d0, b = s.rand(L, 1), s.rand(L, 2), s.rand(L, 3)
    src, dst, rhs = s.fn("src"), s.fn("dst"), s.fn("rhs")
    s.put(src, x, L)
    s.put(dst, d0, L)
    s.put(rhs, b, L)
    s.A.apply(src, dst, L, Inner)
    got = s.get(dst, L)
    want = po.apply(s.o, x, d0, L, Inner)
    m = own(got)
    assert m.sum() < len(got)
    assert pc.rel(got[m], want[m]) < pc.TOL, pc.rel(got[m], want[m])
    # dotGlobal counts every DoF once across the ranks
    mi = s.o.mask(L, Inner)
    assert abs(src.dotGlobal(rhs, L, Inner) - x[mi] @ b[mi]) < 1e-11 * abs(x[mi] @ b[mi])
    assert src.getMaxDoFMagnitude(L, All) == np.abs(x).max()
    cur = x
    for _ in range(3):
        s.A.smooth_jac(dst, rhs, src, 0.66, L, pc.SMOOTH)
        src.assign([1.0], [dst], L, pc.SMOOTH)
        cur = po.smooth_jac(s.o, b, cur, 0.66, L, pc.SMOOTH)
    got = s.get(src, L)
    assert pc.rel(got[m], cur[m]) < pc.TOL
    f = s.fn("f")
    ff, cc = s.rand(L, 10), s.rand(L - 1, 11)
    s.put(f, ff, L)
    s.put(f, cc, L - 1)
    api.P1toP1LinearRestriction().restrict(f, L, All)
    got = s.get(f, L - 1, tol=1e-13)
    want = po.restrict(s.o, ff, cc, L, All)
    mc = own(got)
    assert pc.rel(got[mc], want[mc]) < pc.TOL
    s.put(f, cc, L - 1)
    api.P1toP1LinearProlongation().prolongateAndAdd(f, L - 1, All)
    got = s.get(f, L)
    assert pc.rel(got[m], po.prolongate_and_add(s.o, cc, ff, L - 1, All)[m]) < pc.TOL

    # ---- SOR sweeps and V-cycles: identical to the one-rank run ----------------------------------------------------
    def run(r, w):
        t = pc.Setup(lib, mesh, 0, L, r, w, cell_rank)
        fx, fb, fr = t.fn("x"), t.fn("b"), t.fn("r")
        t.put(fx, x, L)
        t.put(fb, b, L)
        c0, d0 = comm.exchange_stats(lib)
        for k in range(2):
            t.A.smooth_sor(fx, fb, 1.0, L, pc.SMOOTH)
            t.A.smooth_sor(fx, fb, 1.1, L, pc.SMOOTH, backwards=True)
        c1, d1 = comm.exchange_stats(lib)
        if w > 1 and nsend > 0:
            # an exchange carries only the point classes written since the last one (exchange groups): the 4 sweeps make ~ 24
            # exchanges but move little more than the whole ghost volume once per sweep, not once per exchange
            assert c1 - c0 >= 20, (c0, c1)
            assert d1 - d0 <= 6 * nsend, (d1 - d0, nsend, c1 - c0)
        sor = t.get(fx, L)
        fx.interpolate(lambda X, Y, Z: np.sin(X) * np.sinh(Y) + Z, L, DirichletBoundary)
        fx.interpolate(0.0, L, Inner)
        fb.interpolate(0.0, L, All)
        gmg = api.GeometricMultigridSolver(t.st, api.GaussSeidelSmoother(t.st), api.CGSolver(t.st, 0, 0, 100, 1e-16, 1e-16), 0, L, 2, 2)
        res = []
        for _ in range(3):
            gmg.solve(t.A, fx, fb, L)
            t.A.residual(fx, fb, fr, L, Inner)
            res.append(np.sqrt(fr.dotGlobal(fr, L, Inner)))
        if w > 1:
            # the same cycles with the levels 0 .. L-1 on a replica of the whole mesh (no ghost exchange below level L): the same
            # operations on another macro-cell numbering -> the same residual history up to rounding
            c0, _ = comm.exchange_stats(lib)
            fx.interpolate(lambda X, Y, Z: np.sin(X) * np.sinh(Y) + Z, L, DirichletBoundary)
            fx.interpolate(0.0, L, Inner)
            rep = api.ReplicatedCoarseCycle(t.st, L - 1, "gs", 0.66, 2, 2, 100, 1e-16, 1e-16)
            gmr = api.GeometricMultigridSolver(t.st, api.GaussSeidelSmoother(t.st), rep, L - 1, L, 2, 2)
            c1, _ = comm.exchange_stats(lib)
            res_r = []
            for _ in range(3):
                gmr.solve(t.A, fx, fb, L)
                t.A.residual(fx, fb, fr, L, Inner)
                res_r.append(np.sqrt(fr.dotGlobal(fr, L, Inner)))
            c2, _ = comm.exchange_stats(lib)
            assert np.allclose(res_r, res, rtol=1e-9, atol=0), (res_r, res)
            assert c2 - c1 < 40 * 3, (c1, c2)  # only level L exchanges: ~ 30 per cycle instead of ~ 30 per level
            # ... and with level 0 solved directly (cg_max_iter = 0)
            fx.interpolate(lambda X, Y, Z: np.sin(X) * np.sinh(Y) + Z, L, DirichletBoundary)
            fx.interpolate(0.0, L, Inner)
            gmd = api.GeometricMultigridSolver(t.st, api.GaussSeidelSmoother(t.st), api.ReplicatedCoarseCycle(t.st, L - 1, "gs", 0.66, 2, 2, 0), L - 1, L, 2, 2)
            res_d = []
            for _ in range(3):
                gmd.solve(t.A, fx, fb, L)
                t.A.residual(fx, fb, fr, L, Inner)
                res_d.append(np.sqrt(fr.dotGlobal(fr, L, Inner)))
            assert np.allclose(res_d, res, rtol=1e-7, atol=0), (res_d, res)
        return sor, t.get(fx, L), np.array(res)

    sor_p, x_p, res_p = run(rank, world)
    calls, sent = comm.exchange_stats(lib)
    assert calls > 0 and sent > 0
    sor_1, x_1, res_1 = run(0, 1)
    mp = own(sor_p)
    assert np.array_equal(sor_p[mp], sor_1[mp]), np.abs(sor_p[mp] - sor_1[mp]).max()
    assert np.abs(x_p[mp] - x_1[mp]).max() <= 1e-13 * np.abs(x_1).max(), np.abs(x_p[mp] - x_1[mp]).max()
    assert np.allclose(res_p, res_1, rtol=1e-10, atol=0), (res_p, res_1)
    assert res_p[-1] < 0.2 * res_p[0]
    dist.barrier()
    print(f"OK rank {rank}/{world} mesh {mesh_name}: {s.st.num_local_cells} cells, exchange {nsend} doubles, {calls} exchanges, residuals {res_p}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
