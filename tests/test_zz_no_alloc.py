"""hb_p1function_dot_no_alloc: the allocation-free variant of hb_p1function_dot (partial sums in a file-scope __device__ array of the
library instead of a cudaMalloc'ed buffer; include/hyteg_b200.h).  Called through the C-ABI and compared with the existing entry point --
same kernels and summation order, so the results have to agree bitwise (asserted at 1e-13 relative, the BLAS-1 tolerance of the other GPU
tests) -- on two calls in a row, with different data, because every call shares the scratch array and it keeps the last call's contents.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import hyteg_b200 as hb
from hyteg_b200 import api, meshes

pytestmark = pytest.mark.gpu


def _setup(lib, level):
    v, c = meshes.cube_6el()
    st = api.PrimitiveStorage(v, c, lib=lib)
    st.setMeshBoundaryFlagsOnBoundary(1, 0, True)
    a, b = api.P1Function("a", st, level, level), api.P1Function("b", st, level, level)
    return st, a, b


def test_dot_no_alloc_matches_dot_twice(cuda_lib):
    import torch
    lib, level = cuda_lib, 4
    st, a, b = _setup(lib, level)
    rng = np.random.default_rng(7)
    xyz = st.micro_coordinates(level)
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    out_ptr = C.c_void_p(out.data_ptr())
    # the level's tables are built on first use: do that (and resolve the scratch symbol's address) at set-up, outside the checked region
    a.interpolate(1.0, level, hb.All)
    b.interpolate(1.0, level, hb.All)
    a.dotGlobal(b, level, hb.All)
    lib.check(lib.dll.hb_p1function_dot_no_alloc(a.h, b.h, level, hb.All, out_ptr))
    torch.cuda.synchronize()
    for call, flag in enumerate((hb.All, hb.Inner)):
        # consistent copies of the interface DoFs: values given as functions of the coordinates
        w = rng.standard_normal(6)
        a.interpolate(lambda X, Y, Z: w[0] * np.sin(3 * X + w[1]) + w[2] * Y * Z, level, hb.All)
        b.interpolate(lambda X, Y, Z: np.cos(w[3] * X * Y) + w[4] * Z + w[5], level, hb.All)
        expect = a.dotGlobal(b, level, flag)
        free0 = torch.cuda.mem_get_info(0)[0]
        out.fill_(-1.0)
        lib.check(lib.dll.hb_p1function_dot_no_alloc(a.h, b.h, level, flag, out_ptr))
        torch.cuda.synchronize()
        got = float(out.cpu()[0])
        assert torch.cuda.mem_get_info(0)[0] >= free0, "the allocation-free entry point allocated device memory"
        assert abs(expect) > 1e-3
        assert abs(got - expect) <= 1e-13 * abs(expect), (call, got, expect)
    assert len(xyz) == a._n(level)


def test_dot_no_alloc_rejects_what_it_cannot_do(cuda_lib):
    import torch
    lib, level = cuda_lib, 3
    st, a, b = _setup(lib, level)
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    assert lib.dll.hb_p1function_dot_no_alloc(a.h, b.h, level, hb.All, None) != 0
    assert b"null argument" in lib.dll.hb_last_error()
    # boundary flags changed since the level's tables were built: rebuilding them would allocate, so the call refuses
    st.setMeshBoundaryFlagsOnBoundary(2, 0, True)
    assert lib.dll.hb_p1function_dot_no_alloc(a.h, b.h, level, hb.All, C.c_void_p(out.data_ptr())) != 0
    assert b"not built" in lib.dll.hb_last_error()
    a.dotGlobal(b, level, hb.All)  # any ordinary operation rebuilds them
    lib.check(lib.dll.hb_p1function_dot_no_alloc(a.h, b.h, level, hb.All, C.c_void_p(out.data_ptr())))
    torch.cuda.synchronize()
