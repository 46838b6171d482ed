#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY: writes tests/golden/ref_p2.npz -- outputs of the REFERENCE ITSELF for the P2 elementwise path:
P2ElementwiseOperator< forms::p2_div_k_grad_blending_q4 > and < forms::p2_epsilonvar_i_j_blending_q2 > (src/hyteg/elementwiseoperators/
P2ElementwiseOperator.cpp:109-320) with IcosahedralShellMap blending on meshSphericalShell-style meshes, the affine case on cube_6el,
and IcosahedralShellMap::evalF / evalDF samples.  Vectors are stored over oracle/p2_oracle.py's unique-DoF numbering (vertex DoFs, then
edge DoFs); inputs are regenerated from seeds (tests/ref_cases.py: p2_inputs).

Run here (needs /root/reference): python oracle/build_ref.py && python oracle/make_ref_fixtures_p2.py
"""
import os, subprocess, sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hyteg_b200 import meshes
from oracle import p2_oracle as p2
from oracle import ref_bridge as rb

SHELL = (3, 2, 0.5, 1.0)  # the mesh of tests/hyteg/composites/P2P1ElementwiseStokesOperatorTest.cpp:103-155
BARY = np.array([[0.25, 0.25, 0.25, 0.25], [0.7, 0.1, 0.1, 0.1], [0.1, 0.6, 0.2, 0.1], [0.05, 0.15, 0.3, 0.5], [1, 0, 0, 0]])


def p2_inputs(n, seed):
    rng = np.random.default_rng(seed)
    return rng.random(n) - 0.3, rng.random(n)


def main():
    out = {}
    v, c = meshes.spherical_shell(*SHELL)
    s = rb.Session(v, c)
    for L in (2, 3):
        o = p2.P2Oracle(v, c, L, shell=True)
        x, d0 = p2_inputs(o.n, 300 + L)
        out["shell_L%d_divkgrad_inner" % L] = rb.p2_apply(s, o, x, L, form="divkgrad", dst=d0, flag=1, update=0, shell=1)
        if L == 2:
            out["shell_L2_divkgrad_add_all"] = rb.p2_apply(s, o, x, L, form="divkgrad", dst=d0, flag=15, update=1, shell=1)
            for i, j in ((0, 0), (0, 1), (1, 2), (2, 0), (2, 2)):
                out["shell_L2_epsilon_%d%d_inner" % (i, j)] = rb.p2_apply(s, o, x, L, form="epsilon", dst=d0, flag=1, update=0, shell=1, i=i, j=j)
        print("shell level", L, "done", flush=True)
    # B and B^T blocks of the Stokes operator on the shell (P2ToP1ElementwiseBlendingDiv*, P1ToP2ElementwiseBlendingDivT*), level 2
    o = p2.P2Oracle(v, c, 2, shell=True)
    x, d0 = p2_inputs(o.n, 302)
    phys = o.dof_physical_coordinates()
    tol = 0.05 * rb.min_spacing(v, c, 2)
    for comp in (0, 1, 2):
        oute = os.path.join(s.dir, "oe.bin")
        _, cl = s.run("p2top1", 2, src=s.cloud(phys, x), oute=oute, comp=comp, flag=1, shell=1)
        out["shell_L2_div_%d_inner" % comp] = rb.to_global(o.xyz[: o.nv], cl, tol)
        _, cl = s.run("p2fromp1", 2, src=s.cloud(phys[: o.nv], d0[: o.nv]), oute=oute, comp=comp, flag=1, shell=1)
        de = np.fromfile(oute).reshape(-1, 4)
        out["shell_L2_divt_%d_inner" % comp] = np.concatenate([rb.to_global(o.xyz[: o.nv], cl, tol), rb.to_global(o.xyz[o.nv:], (de[:, :3], de[:, 3]), tol)])
    # Chebyshev smoother and the Chebyshev-smoothed GMG on the blended div-k-grad operator (levels 1..3, V(2,2), order 3, CG on level 1)
    o = p2.P2Oracle(v, c, 3, shell=True)
    rng = np.random.default_rng(9)
    x0, b, ev = rng.random(o.n), rng.random(o.n) - 0.5, rng.random(o.n)
    phys = o.dof_physical_coordinates()
    cx, cb, cev = s.cloud(phys, x0), s.cloud(phys, b), s.cloud(phys, ev)
    tol = 0.05 * rb.min_spacing(v, c, 3)

    def getp2(cl, oute):
        de = np.fromfile(oute).reshape(-1, 4)
        return np.concatenate([rb.to_global(o.xyz[: o.nv], cl, tol), rb.to_global(o.xyz[o.nv:], (de[:, :3], de[:, 3]), tol)])

    oute = os.path.join(s.dir, "ge.bin")
    r, cl = s.run("p2cheb", 3, minlevel=3, x=cx, rhs=cb, ev=cev, order=3, sweeps=2, shell=1, oute=oute)
    out["shell_L3_cheb3_2sweeps"] = getp2(cl, oute)
    out["shell_L3_rho20"] = np.array(r["rho"])
    r, cl = s.run("p2gmg", 3, minlevel=1, x=cx, rhs=cb, ev=cev, order=3, cycles=5, pre=2, post=2, shell=1, oute=oute)
    out["shell_gmg_cheb_res2"] = np.array(r["res2"])
    out["shell_gmg_cheb_x"] = getp2(cl, oute)
    print("P2 GMG done", flush=True)
    # IcosahedralShellMap samples per macro-cell (cells identified by their centroid)
    for nt, nr in ((3, 2), (2, 3)):
        vv, cc = meshes.spherical_shell(nt, nr, 0.5, 1.0)
        ss = rb.Session(vv, cc)
        f = os.path.join(ss.dir, "sm.bin")
        subprocess.check_call([rb.REF_DUMP, "shellmap", "mesh=" + ss.msh, "shell=1", "out=" + f], stdout=subprocess.DEVNULL)
        out["shellmap_%d_%d" % (nt, nr)] = np.fromfile(f).reshape(len(cc), 72)
    v, c = meshes.cube_6el()
    s = rb.Session(v, c)
    for L in (2, 3):
        o = p2.P2Oracle(v, c, L, shell=False)
        x, d0 = p2_inputs(o.n, 300 + L)
        out["cube_6el_L%d_divkgrad_inner" % L] = rb.p2_apply(s, o, x, L, form="divkgrad", dst=d0, flag=1, update=0, shell=0)
    # P2toP2QuadraticProlongation / Restriction (flag Inner | NeumannBoundary = 5): coarse level Lc <-> Lc + 1
    for kind, Lc in (("cube", 2), ("shell", 1)):
        v, c = meshes.cube_6el() if kind == "cube" else meshes.spherical_shell(*SHELL)
        s = rb.Session(v, c)
        oc, of = p2.P2Oracle(v, c, Lc, shell=False), p2.P2Oracle(v, c, Lc + 1, shell=False)
        xc, _ = p2_inputs(oc.n, 400 + Lc)
        xf, _ = p2_inputs(of.n, 500 + Lc)

        def run(cmd, L, o_src, x, o_dst):
            oute = os.path.join(s.dir, "te.bin")
            _, cl = s.run(cmd, L, src=s.cloud(o_src.xyz, x), oute=oute, flag=5)
            de = np.fromfile(oute).reshape(-1, 4)
            tol = 0.05 * rb.min_spacing(v, c, o_dst.L)
            return np.concatenate([rb.to_global(o_dst.xyz[: o_dst.nv], cl, tol), rb.to_global(o_dst.xyz[o_dst.nv:], (de[:, :3], de[:, 3]), tol)])

        out["%s_Lc%d_prolongate" % (kind, Lc)] = run("p2prolongate", Lc, oc, xc, of)
        out["%s_Lc%d_restrict" % (kind, Lc)] = run("p2restrict", Lc + 1, of, xf, oc)
    p = os.path.join(ROOT, "tests", "golden", "ref_p2.npz")
    np.savez_compressed(p, **out)
    print("wrote", p, os.path.getsize(p) // 1024, "KiB,", len(out), "arrays")


if __name__ == "__main__":
    main()
