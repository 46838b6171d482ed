"""BASELINE config 3: P2 elementwise variable-coefficient div-k-grad with IcosahedralShellMap blending on the thick spherical shell
(meshSphericalShell(3,2,0.5,1.0), 240 macro-cells); CUDA-event times of one apply per level."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hyteg_b200 as hb
from hyteg_b200 import api, meshes
levels = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [4, 5, 6]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
lib = hb.get_lib(); lib.check(lib.dll.hb_device_init(0))
s = torch.cuda.current_stream(); lib.dll.hb_set_stream(s.cuda_stream)
v, c = meshes.spherical_shell(3, 2, 0.5, 1.0)
def t(fn, n):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s)
    for _ in range(n): fn()
    b.record(s); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for L in levels:
    st = api.PrimitiveStorage(v, c); st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    api.setIcosahedralShellMap(st)
    N = 1 << L
    src, dst, k = (api.P2Function(n, st, L, L) for n in ("src", "dst", "k"))
    rng = np.random.default_rng(0)
    nv, ne = src.v._n(L), src.edge_size(L)
    src.upload(rng.random(nv), rng.random(ne), L)   # synthetic data; copies of interface DoFs need not agree for a timing run
    k.interpolate(1.5, L, hb.All)
    A = api.P2ElementwiseBlendingDivKGradOperator(st, L, L, k)
    us = t(lambda: A.apply(src, dst, L, hb.Inner), reps)
    n_el = len(c) * 8 ** L
    # unique DoFs ~ (vertices + edges) of the refined mesh: count entries whose copy is the owner is not exposed; use element count based estimate
    ndof = st.num_global_dofs(L) + 0  # vertex DoFs
    nedge = int(round(n_el * 6 / 5.0))  # ~1.2 edges per micro-tet asymptotically (each interior edge shared by ~5 tets)
    print(json.dumps({"level": L, "micro_tets": n_el, "us_per_apply": round(us, 1), "Gtets_per_s": round(n_el / us * 1e-3, 3),
                      "vertex_dofs": ndof, "approx_GDoF_per_s": round((ndof + nedge) / us * 1e-3, 3)}), flush=True)
    # BASELINE config 4: P2-P1 Taylor-Hood Stokes operator with the epsilon velocity block (fused vector kernel + B + B^T)
    S = api.P2P1ElementwiseBlendingEpsilonStokesOperator(st, L, L, k)
    ts, td = api.P2P1TaylorHoodFunction("s", st, L, L), api.P2P1TaylorHoodFunction("d", st, L, L)
    for f in ts.uvw:
        f.upload(rng.random(nv), rng.random(ne), L)
    ts.p.upload(rng.random(nv), L)
    us_eps = t(lambda: S.A.apply(ts.uvw, td.uvw, L, hb.Inner), reps)
    us_st = t(lambda: S.apply(ts, td, L, hb.Inner), reps)
    fused, S.fused = S.fused, None
    us_unf = t(lambda: S.apply(ts, td, L, hb.Inner), reps)
    S.fused = fused
    print(json.dumps({"level": L, "epsilon_vector_us": round(us_eps, 1), "stokes_apply_us": round(us_st, 1), "stokes_apply_7_kernels_us": round(us_unf, 1),
                      "Gtets_per_s_epsilon": round(n_el / us_eps * 1e-3, 3), "stokes_dofs": 3 * (ndof + nedge) + ndof,
                      "stokes_GDoF_per_s": round((3 * (ndof + nedge) + ndof) / us_st * 1e-3, 3)}), flush=True)
    del A, src, dst, k, st, S, ts, td
