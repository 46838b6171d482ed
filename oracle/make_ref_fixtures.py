#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY: writes tests/golden/ref_p1.npz -- outputs of the REFERENCE ITSELF (HyTeG's
P1ConstantLaplaceOperator, smoothers, transfer operators, CGSolver, GeometricMultigridSolver; built from /root/reference by
oracle/build_ref.py, driven by oracle/ref_dump.cpp) on seeded inputs.  tests/test_oracle_pinning.py pins the oracle to them and
the -m gpu tests compare the CUDA path with them.  Vectors are stored over the oracle's unique-DoF numbering
(P1Oracle.levels[L].xyz order), inputs are regenerated from the stored seed.

Run here (needs /root/reference): python oracle/build_ref.py && python oracle/make_ref_fixtures.py
"""
import os, sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hyteg_b200 import meshes
from oracle import p1_oracle as po
from oracle import ref_bridge as rb


def inputs(o, L, seed):
    """seeded test vectors over the unique DoFs of level L (shared with the tests: tests/ref_cases.py imports this)"""
    rng = np.random.default_rng(seed)
    lv = o.levels[L]
    x = rng.random(lv.n)
    b = rng.random(lv.n) - 0.5
    return x, b


def main():
    out = {}
    for name, levels in (("cube_6el", (3, 4, 5)), ("three_tets_with_two_joint_faces", (3, 4)), ("regular_octahedron_8el", (3, 4))):
        v, c = meshes.golden(name)
        o = po.P1Oracle(v, c, 0, max(levels))
        s = rb.Session(v, c)
        for L in levels:
            lv = o.levels[L]
            tol = 0.1 * rb.min_spacing(v, c, L)
            x, b = inputs(o, L, 100 + L)
            cx, cb = s.cloud(lv.xyz, x), s.cloud(lv.xyz, b)
            G = lambda cl: rb.to_global(lv.xyz, cl, tol)
            k = "%s_L%d_" % (name, L)
            # Operator::apply, Replace on Inner and Add on All
            out[k + "apply_inner"] = G(s.run("apply", L, src=cx, dst=cb, flag=1)[1])
            out[k + "apply_add_all"] = G(s.run("apply", L, src=cx, dst=cb, flag=15, update=1)[1])
            # smooth_sor forward / backward (one sweep, omega 1 and 1.3), smooth_jac, SGS
            out[k + "sor_fwd"] = G(s.run("sor", L, x=cx, rhs=cb, omega=1.0)[1])
            out[k + "sor_bwd"] = G(s.run("sor", L, x=cx, rhs=cb, omega=1.0, backwards=1)[1])
            if L == 5:
                continue  # the largest level: apply and one SOR sweep each way only (keeps the fixture file small)
            out[k + "sor_fwd_w13_3"] = G(s.run("sor", L, x=cx, rhs=cb, omega=1.3, sweeps=3)[1])
            out[k + "jac_3"] = G(s.run("jac", L, x=cx, rhs=cb, omega=2.0 / 3.0, sweeps=3)[1])
            # Chebyshev order 3 and 5 with the radius of 20 power iterations started from x
            for order in (3, 5):
                r, cl = s.run("cheb", L, x=cx, rhs=cb, ev=cx, order=order, power_iterations=20)
                out[k + "cheb%d" % order] = G(cl)
                out[k + "rho20"] = np.array(r["rho"])
            # restriction of level L -> L-1 and prolongation L-1 -> L
            lc = o.levels[L - 1]
            tolc = 0.1 * rb.min_spacing(v, c, L - 1)
            out[k + "restrict"] = rb.to_global(lc.xyz, s.run("restrict", L, src=cx, flag=5)[1], tolc)
            xc, _ = inputs(o, L - 1, 200 + L)
            ccx = s.cloud(lc.xyz, xc)
            out[k + "prolongate"] = G(s.run("prolongate", L - 1, src=ccx, flag=5)[1])
        # V(3,3) cycles with each smoother, levels 0..max: squared-residual history + final iterate
        L = max(levels)
        lv = o.levels[L]
        tol = 0.1 * rb.min_spacing(v, c, L)
        x, b = inputs(o, L, 100 + L)
        # P1GMG3DConvergenceTest.cpp:96-101: random inner values, boundary values of a smooth function, f = 0
        m = o.mask(L, po.INNER)
        u0 = np.sin(lv.xyz[:, 0]) * np.sinh(lv.xyz[:, 1]) * lv.xyz[:, 2]
        u0[m] = x[m]
        cu, cz = s.cloud(lv.xyz, u0), s.cloud(lv.xyz, np.zeros(lv.n))
        for sm, kv in (("gs", {}), ("sor", {"omega": 1.2}), ("sgs", {}), ("jac", {"omega": 2.0 / 3.0}), ("cheb", {"order": 3, "ev": cu})):
            r, cl = s.run("gmg", L, minlevel=0, x=cu, rhs=cz, cycles=10, smoother=sm, **kv)
            out["%s_gmg_%s_res2" % (name, sm)] = np.array(r["res2"])
            out["%s_gmg_%s_x" % (name, sm)] = rb.to_global(lv.xyz, cl, tol)
        # CG to 1e-16 (P1CG3DConvergenceTest.cpp:103-146 on the octahedron, levels 3 and 4)
        for Lc in levels[-2:]:
            lvc = o.levels[Lc]
            xx, _ = inputs(o, Lc, 100 + Lc)
            mm = o.mask(Lc, po.INNER)
            uu = np.sin(lvc.xyz[:, 0]) * np.sinh(lvc.xyz[:, 1]) * lvc.xyz[:, 2]
            uex = uu.copy()
            uu[mm] = xx[mm]
            r, cl = s.run("cg", Lc, x=s.cloud(lvc.xyz, uu), rhs=s.cloud(lvc.xyz, np.zeros(lvc.n)))
            xs = rb.to_global(lvc.xyz, cl, 0.1 * rb.min_spacing(v, c, Lc))
            out["%s_L%d_cg_x" % (name, Lc)] = xs
            out["%s_L%d_cg_res2" % (name, Lc)] = np.array(r["res2"])
            out["%s_L%d_cg_err" % (name, Lc)] = np.array([np.sqrt(((xs - uex)[mm] ** 2).sum() / mm.sum())])
        if name == "cube_6el":
            # ChebyshevSmoother::setupCoefficientsInternal, orders 1..10 (closed forms, ChebyshevSmoother.hpp:341-539)
            r, _ = s.run("chebcoeffs", 2, lower=0.9, upper=3.7)
            for order in range(1, 11):
                out["cheb_coeffs_order%d" % order] = np.array(r["order%d" % order])
        print(name, "done", flush=True)
    p = os.path.join(ROOT, "tests", "golden", "ref_p1.npz")
    np.savez_compressed(p, **out)
    print("wrote", p, os.path.getsize(p) // 1024, "KiB,", len(out), "arrays")


if __name__ == "__main__":
    main()
