"""CPU-only tests of the host logic (topology tables, orientation maps, solver drivers) through the
C-ABI of the TEST-ONLY emulation build.  The product path is covered by tests/test_gpu_parity.py."""
import numpy as np
import pytest

from hyteg_b200 import meshes
from hyteg_b200.capi import Inner, All, Add, Replace, DirichletBoundary
import parity_cases as pc

MESHES = ["cube_6el", "regular_octahedron_8el", "pyramid_tilted_4el", "three_tets_with_two_joint_faces", "tet_tilted_1el", "cube_24el"]


@pytest.mark.parametrize("name", MESHES)
@pytest.mark.parametrize("L", [0, 1, 2, 3])
def test_apply(emul_lib, name, L):
    pc.case_apply(emul_lib, meshes.golden(name), L, Inner)
    pc.case_apply(emul_lib, meshes.golden(name), L, All, Add)


@pytest.mark.parametrize("name", ["cube_6el", "regular_octahedron_8el"])
def test_apply_level4(emul_lib, name):
    pc.case_apply(emul_lib, meshes.golden(name), 4, Inner | DirichletBoundary)


@pytest.mark.parametrize("name", MESHES[:4])
@pytest.mark.parametrize("L", [2, 3])
def test_linear_functions_in_kernel(emul_lib, name, L):
    pc.case_linear_kernel(emul_lib, meshes.golden(name), L)


@pytest.mark.parametrize("L", [1, 3])
def test_jacobi(emul_lib, L):
    pc.case_jacobi(emul_lib, meshes.golden("cube_6el"), L)


@pytest.mark.parametrize("order", [1, 2, 3, 4])
def test_chebyshev(emul_lib, order):
    pc.case_chebyshev(emul_lib, meshes.golden("regular_octahedron_8el"), 3, order)


@pytest.mark.parametrize("name", ["cube_6el", "pyramid_tilted_4el"])
def test_blas(emul_lib, name):
    pc.case_blas(emul_lib, meshes.golden(name), 3)


@pytest.mark.parametrize("name", ["cube_6el", "regular_octahedron_8el", "three_tets_with_two_joint_faces"])
@pytest.mark.parametrize("Lc", [0, 1, 2])
def test_transfer(emul_lib, name, Lc):
    pc.case_transfer(emul_lib, meshes.golden(name), Lc)


@pytest.mark.parametrize("name,n_ranks", [("cube_6el", 2), ("cube_24el", 3), ("regular_octahedron_8el", 4)])
@pytest.mark.parametrize("L", [1, 3])
def test_exchange_plan_is_symmetric(emul_lib, name, n_ranks, L):
    """ghost exchange plan: what all ranks send equals what all ranks receive; every rank with a neighbour exchanges something;
    the mirrored layers are a small fraction of a macro-cell on finer levels"""
    from hyteg_b200 import api
    v, c = meshes.golden(name)
    sent = recv = 0
    for r in range(n_ranks):
        st = api.PrimitiveStorage(v, c, r, n_ranks, lib=emul_lib)
        st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
        peers, ns, nr = st.exchange_info(L)
        assert 1 <= peers <= n_ranks - 1
        assert ns > 0 and nr > 0
        sent += ns
        recv += nr
    assert sent == recv


def test_interface_coordinates_identical_in_all_copies(emul_lib):
    """micro-vertex coordinates of an interface point are bit-identical in every macro-cell that holds a copy, also on a mesh
    with irrational vertex coordinates (spherical shell): interpolate( expression ) keeps the copies consistent"""
    from hyteg_b200 import api
    v, c = meshes.spherical_shell(2, 2)
    L = 2
    st = api.PrimitiveStorage(v, c, lib=emul_lib)
    st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    f = api.P1Function("f", st, L, L)
    f.interpolate(lambda x, y, z: np.sin(3 * x) * np.exp(y) + z * z, L, All)
    from oracle import p1_oracle as po
    o = po.P1Oracle(v, c, L, L)
    o.from_cells(f.download(L), L, check=True, tol=0.0)  # raises if two copies of a DoF differ


@pytest.mark.parametrize("mesh", ["shell23", "cuboid322"])
def test_direct_level0_solver_matches_cg(emul_lib, mesh):
    """dense Cholesky of the level-0 operator on the free macro vertices == CG driven to round-off, also with non-zero Dirichlet data, and as
    the coarse-grid solver of a V-cycle"""
    from hyteg_b200 import api
    import hyteg_b200 as hb
    v, c = meshes.spherical_shell(2, 4, 0.5, 1.0) if mesh == "shell23" else meshes.cuboid(3, 2, 2)
    st = api.PrimitiveStorage(v, c, lib=emul_lib)
    st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    L = 2
    A = api.P1ConstantLaplaceOperator(st, 0, L)
    x, y, b = (api.P1Function(n, st, 0, L) for n in "xyb")
    g = lambda X, Y, Z: np.sin(2 * X) + Y * Z
    for f in (x, y):
        f.interpolate(0.0, 0, All)
        f.interpolate(g, 0, DirichletBoundary)
    b.interpolate(lambda X, Y, Z: 1.0 + X - Z, 0, All)
    assert st.num_global_dofs(0) > 20
    api.CGSolver(st, 0, 0, 500, 1e-30, 1e-13).solve(A, x, b, 0)
    api.DirectLevel0Solver(st).solve(A, y, b, 0)
    xa, ya = x.download(0), y.download(0)
    assert np.abs(xa).max() > 0.1
    assert np.abs(xa - ya).max() < 1e-10 * np.abs(xa).max(), np.abs(xa - ya).max()
    # inside a V-cycle: same residual history as with CG on level 0
    hist = []
    for coarse in (api.CGSolver(st, 0, 0, 500, 1e-30, 1e-14), api.DirectLevel0Solver(st)):
        gmg = api.GeometricMultigridSolver(st, api.GaussSeidelSmoother(st), coarse, 0, L, 2, 2)
        x.interpolate(0.0, L, All)
        x.interpolate(g, L, DirichletBoundary)
        b.interpolate(0.0, L, All)
        r = []
        for _ in range(3):
            gmg.solve(A, x, b, L)
            A.residual(x, b, y, L, Inner)
            r.append(np.sqrt(y.dotGlobal(y, L, Inner)))
        hist.append(r)
    assert np.allclose(hist[0], hist[1], rtol=1e-8, atol=0), hist
    assert hist[0][-1] < 0.1 * hist[0][0]
