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
