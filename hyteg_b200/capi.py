"""ctypes binding of the hyteg_b200 C-ABI (include/hyteg_b200.h).

The package only ever loads the CUDA library ``hyteg_b200/libhyteg_b200.so`` (built by
``__graft_entry__.build()``); there is no CPU fallback -- a missing library or a missing device
raises.  ``Lib(path)`` accepts an explicit path so that the unit tests can bind the test-only host
emulation build of the same sources (tests/emul/), which is never used by the product path.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(HERE, "libhyteg_b200.so")

# DoFType / UpdateType (reference src/hyteg/types/types.hpp:29-44)
Inner, DirichletBoundary, NeumannBoundary, FreeslipBoundary = 1, 2, 4, 8
Boundary, All, NoneFlag = 14, 15, 0
Replace, Add = 0, 1
VCYCLE, WCYCLE = 0, 1

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_vp = C.c_void_p

_SIGNATURES = {
    "hb_last_error": (C.c_char_p, []),
    "hb_version": (C.c_char_p, []),
    "hb_device_init": (C.c_int, [C.c_int]),
    "hb_set_stream": (C.c_int, [_vp]),
    "hb_synchronize": (C.c_int, []),
    "hb_set_allreduce": (C.c_int, [_vp, _vp]),
    "hb_set_tuning": (C.c_int, [C.c_char_p, C.c_int]),
    "hb_nccl_unique_id": (C.c_int, [_vp]),
    "hb_nccl_init": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_nccl_finalize": (C.c_int, []),
    "hb_p2p_window": (C.c_int, [_vp, C.c_longlong, C.c_int]),
    "hb_p2p_connect": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_p2p_finalize": (C.c_int, []),
    "hb_p2p_active": (C.c_int, [C.POINTER(C.c_longlong)]),
    "hb_set_exchange": (C.c_int, [_vp, _vp]),
    "hb_exchange_stats": (C.c_int, [C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "hb_storage_exchange_info": (C.c_int, [_vp, C.c_int, _ip, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]),
    "hb_p1function_communicate": (C.c_int, [_vp, C.c_int]),
    "hb_storage_create": (_vp, [C.c_int, _dp, C.c_int, _ip, C.c_int, C.c_int]),
    "hb_storage_create_partitioned": (_vp, [C.c_int, _dp, C.c_int, _ip, C.c_int, C.c_int, _ip]),
    "hb_storage_destroy": (None, [_vp]),
    "hb_storage_set_mesh_boundary_flags_on_boundary": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "hb_storage_counts": (C.c_int, [_vp, _ip]),
    "hb_storage_primitive_vertices": (C.c_int, [_vp, C.c_int, _ip]),
    "hb_storage_get_mesh_boundary_flags": (C.c_int, [_vp, C.c_int, _ip]),
    "hb_storage_set_mesh_boundary_flags": (C.c_int, [_vp, C.c_int, _ip]),
    "hb_storage_num_local_cells": (C.c_int, [_vp]),
    "hb_storage_local_cell_global_id": (C.c_int, [_vp, C.c_int]),
    "hb_cell_array_size": (C.c_longlong, [C.c_int]),
    "hb_storage_num_global_dofs": (C.c_longlong, [_vp, C.c_int]),
    "hb_storage_micro_coordinates": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p1function_create": (_vp, [_vp, C.c_int, C.c_int]),
    "hb_p1function_destroy": (None, [_vp]),
    "hb_p1function_set_boundary_condition": (C.c_int, [_vp, _ip, C.c_int]),
    "hb_p1function_copy_boundary_condition": (C.c_int, [_vp, _vp]),
    "hb_p1function_storage": (_vp, [_vp]),
    "hb_p1function_level_range": (C.c_int, [_vp, _ip, _ip]),
    "hb_p1function_upload": (C.c_int, [_vp, C.c_int, _dp, C.c_int]),
    "hb_p1function_download": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p1function_upload_async": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p1function_download_async": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p1function_upload_cell": (C.c_int, [_vp, C.c_int, C.c_int, _dp]),
    "hb_p1function_download_cell": (C.c_int, [_vp, C.c_int, C.c_int, _dp]),
    "hb_p1function_device_ptr": (_vp, [_vp, C.c_int]),
    "hb_p1function_interpolate_constant": (C.c_int, [_vp, C.c_double, C.c_int, C.c_int]),
    "hb_p1function_assign": (C.c_int, [_vp, C.c_int, _dp, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p1function_add": (C.c_int, [_vp, C.c_int, _dp, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p1function_add_scalar": (C.c_int, [_vp, C.c_double, C.c_int, C.c_int]),
    "hb_p1function_mult_elementwise": (C.c_int, [_vp, C.c_int, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p1function_dot": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _dp]),
    "hb_storage_create_replica": (_vp, [_vp]),
    "hb_p1function_get_boundary_condition": (C.c_int, [_vp, C.POINTER(C.c_int), C.c_int]),
    "hb_solver_create_direct_level0": (_vp, [_vp]),
    "hb_p1function_gather_to_replica": (C.c_int, [_vp, _vp, C.c_int]),
    "hb_p1function_scatter_from_replica": (C.c_int, [_vp, _vp, C.c_int]),
    "hb_solver_create_replicated_coarse_cycle": (_vp, [_vp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]),
    "hb_p1function_dot_no_alloc": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _vp]),  # out: DEVICE pointer
    "hb_p1function_sum": (C.c_int, [_vp, C.c_int, C.c_int, _dp]),
    "hb_p1function_max_magnitude": (C.c_int, [_vp, C.c_int, C.c_int, _dp]),
    "hb_p1_constant_laplace_create": (_vp, [_vp, C.c_int, C.c_int]),
    "hb_operator_destroy": (None, [_vp]),
    "hb_operator_level_range": (C.c_int, [_vp, _ip, _ip]),
    "hb_operator_apply": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int]),
    "hb_operator_smooth_jac": (C.c_int, [_vp, _vp, _vp, _vp, C.c_double, C.c_int, C.c_int]),
    "hb_operator_smooth_sor": (C.c_int, [_vp, _vp, _vp, C.c_double, C.c_int, C.c_int, C.c_int]),
    "hb_operator_smooth_sor_ordered": (C.c_int, [_vp, _vp, _vp, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int]),
    "hb_operator_residual": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int]),
    "hb_operator_cheb_first": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int]),
    "hb_operator_cheb_step": (C.c_int, [_vp, _vp, _vp, _vp, C.c_double, C.c_int, C.c_int]),
    "hb_operator_diagonal": (C.c_int, [_vp, _vp, C.c_int, C.c_int]),
    "hb_operator_cell_stencil": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _dp]),
    "hb_p1_restrict": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_p1_prolongate": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_p1_prolongate_and_add": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_storage_set_icosahedral_shell_map": (C.c_int, [_vp]),
    "hb_storage_shell_map_parameters": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_storage_map_eval": (C.c_int, [_vp, C.c_int, C.c_int, _dp, C.c_int, _dp]),
    "hb_p2function_create": (_vp, [_vp, C.c_int, C.c_int]),
    "hb_p2function_destroy": (None, [_vp]),
    "hb_p2function_vertex_part": (_vp, [_vp]),
    "hb_p2_edge_cell_size": (C.c_longlong, [C.c_int]),
    "hb_p2function_upload_edges": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p2function_download_edges": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p2function_edge_device_ptr": (_vp, [_vp, C.c_int]),
    "hb_p2_edge_endpoints": (C.c_int, [_vp, C.c_int, C.c_int, _dp]),
    "hb_p2function_interpolate_constant": (C.c_int, [_vp, C.c_double, C.c_int, C.c_int]),
    "hb_p2function_assign": (C.c_int, [_vp, C.c_int, _dp, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p2function_add": (C.c_int, [_vp, C.c_int, _dp, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p2function_dot": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _dp]),
    "hb_p2_divkgrad_create": (_vp, [_vp, C.c_int, C.c_int, _vp]),
    "hb_p2_epsilon_create": (_vp, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp]),
    "hb_p2operator_destroy": (None, [_vp]),
    "hb_p2operator_num_quadrature_points": (C.c_longlong, [_vp, C.c_int]),
    "hb_p2operator_quadrature_points": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p2operator_set_sampled_coefficient": (C.c_int, [_vp, C.c_int, _dp]),
    "hb_p2operator_apply": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int]),
    "hb_p2function_mult_elementwise": (C.c_int, [_vp, C.c_int, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p2function_copy_boundary_condition": (C.c_int, [_vp, _vp]),
    "hb_p2operator_diagonal": (C.c_int, [_vp, C.POINTER(_vp), C.c_int, C.c_int]),
    "hb_p2_gmg_chebyshev_create": (_vp, [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double]),
    "hb_p2solver_spectral_radius": (C.c_double, [_vp]),
    "hb_p2solver_solve": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int]),
    "hb_p2solver_destroy": (None, [_vp]),
    "hb_p2_stokes_create": (_vp, [_vp, C.c_int, C.c_int, _vp]),
    "hb_p2operator_apply_stokes": (C.c_int, [_vp, C.POINTER(_vp), _vp, C.POINTER(_vp), _vp, C.c_int, C.c_int]),
    "hb_p2_div_create": (_vp, [_vp, C.c_int, C.c_int, C.c_int]),
    "hb_p2_divt_create": (_vp, [_vp, C.c_int, C.c_int, C.c_int]),
    "hb_p2_epsilon_vector_create": (_vp, [_vp, C.c_int, C.c_int, _vp]),
    "hb_p2operator_apply_p2_to_p1": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int]),
    "hb_p2operator_apply_p1_to_p2": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int]),
    "hb_p2operator_apply_vector": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.c_int, C.c_int, C.c_int]),
    "hb_p2_restrict": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_p2_prolongate": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_p2_prolongate_and_add": (C.c_int, [_vp, C.c_int, C.c_int]),
    "hb_solver_create_cg": (_vp, [_vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]),
    "hb_solver_create_weighted_jacobi": (_vp, [_vp, C.c_int, C.c_int, C.c_double]),
    "hb_solver_create_sor": (_vp, [C.c_double]),
    "hb_solver_create_gauss_seidel": (_vp, []),
    "hb_solver_create_symmetric_gauss_seidel": (_vp, []),
    "hb_solver_create_chebyshev": (_vp, [_vp, C.c_int, C.c_int]),
    "hb_chebyshev_setup_coefficients": (C.c_int, [_vp, C.c_int, C.c_double, C.c_double]),
    "hb_chebyshev_coefficients": (C.c_int, [_vp, _dp, _ip]),
    "hb_estimate_spectral_radius": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, _dp]),
    "hb_solver_create_gmg": (_vp, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "hb_solver_solve": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int]),
    "hb_solver_iterations": (C.c_int, [_vp]),
    "hb_solver_destroy": (None, [_vp]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES.keys())


class HytegB200Error(RuntimeError):
    pass


class Lib:
    """A loaded C-ABI library with typed entry points."""

    def __init__(self, path=DEFAULT_LIB):
        if not os.path.exists(path):
            raise HytegB200Error(
                f"hyteg_b200: native library {path} not found -- run __graft_entry__.build(); "
                "there is no CPU fallback"
            )
        self.path = path
        self.dll = C.CDLL(path, mode=C.RTLD_GLOBAL)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(self.dll, name)
            fn.restype = res
            fn.argtypes = args
        self._keepalive = []

    def check(self, rc):
        if rc != 0:
            raise HytegB200Error(self.dll.hb_last_error().decode())

    def checkp(self, p):
        if not p:
            raise HytegB200Error(self.dll.hb_last_error().decode())
        return p

    def version(self):
        return self.dll.hb_version().decode()


_default = None


def get_lib():
    """The CUDA library of the package (loaded once)."""
    global _default
    if _default is None:
        _default = Lib(DEFAULT_LIB)
    return _default


def as_dp(a):
    return a.ctypes.data_as(_dp)


def as_ip(a):
    return a.ctypes.data_as(_ip)


def f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)
