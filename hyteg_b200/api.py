"""Python mirror of the HyTeG interfaces of the P1 multigrid path, over the C-ABI.

Class and method names follow the reference (P1Function, P1ConstantLaplaceOperator,
GeometricMultigridSolver, ...; see include/hyteg_b200/hyteg.hpp for the file:line citations) so the
parity tests read like HyTeG's own tests.  All arithmetic runs in the CUDA library.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (Add, All, Boundary, DirichletBoundary, FreeslipBoundary, Inner, NeumannBoundary,
                   Replace, VCYCLE, WCYCLE, as_dp, as_ip, f64, i32)

SMOOTH_FLAG = Inner | NeumannBoundary | FreeslipBoundary


class PrimitiveStorage:
    """SetupPrimitiveStorage + PrimitiveStorage for tetrahedral macro meshes."""

    def __init__(self, vertices, cells, rank=0, n_ranks=1, lib=None, cell_rank=None):
        """cell_rank: target rank of every macro-cell (e.g. meshes.partition_greedy); None: contiguous blocks"""
        self.lib = lib or capi.get_lib()
        self.vertices = f64(vertices).reshape(-1, 3)
        self.cells = i32(cells).reshape(-1, 4)
        if cell_rank is None:
            self.h = self.lib.checkp(self.lib.dll.hb_storage_create(
                len(self.vertices), as_dp(self.vertices), len(self.cells), as_ip(self.cells), rank, n_ranks))
        else:
            cr = i32(cell_rank)
            assert len(cr) == len(self.cells)
            self.h = self.lib.checkp(self.lib.dll.hb_storage_create_partitioned(
                len(self.vertices), as_dp(self.vertices), len(self.cells), as_ip(self.cells), rank, n_ranks, as_ip(cr)))
        self.rank, self.n_ranks = rank, n_ranks

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_storage_destroy(self.h)
            self.h = None

    def setMeshBoundaryFlagsOnBoundary(self, flag_on_boundary, flag_inner, highest_dim_always_inner):
        self.lib.check(self.lib.dll.hb_storage_set_mesh_boundary_flags_on_boundary(
            self.h, flag_on_boundary, flag_inner, int(highest_dim_always_inner)))

    def counts(self):
        c = np.zeros(4, dtype=np.int32)
        self.lib.check(self.lib.dll.hb_storage_counts(self.h, as_ip(c)))
        return tuple(int(x) for x in c)

    def primitive_vertices(self, dim):
        n = self.counts()[dim]
        out = np.zeros((n, dim + 1), dtype=np.int32)
        self.lib.check(self.lib.dll.hb_storage_primitive_vertices(self.h, dim, as_ip(out)))
        return out

    def get_mesh_boundary_flags(self, dim):
        out = np.zeros(self.counts()[dim], dtype=np.int32)
        self.lib.check(self.lib.dll.hb_storage_get_mesh_boundary_flags(self.h, dim, as_ip(out)))
        return out

    def set_mesh_boundary_flags(self, dim, flags):
        flags = i32(flags)
        assert len(flags) == self.counts()[dim]
        self.lib.check(self.lib.dll.hb_storage_set_mesh_boundary_flags(self.h, dim, as_ip(flags)))

    @property
    def num_local_cells(self):
        return self.lib.dll.hb_storage_num_local_cells(self.h)

    def local_cell_global_ids(self):
        return [self.lib.dll.hb_storage_local_cell_global_id(self.h, i) for i in range(self.num_local_cells)]

    def cell_array_size(self, level):
        return int(self.lib.dll.hb_cell_array_size(level))

    def num_global_dofs(self, level):
        return int(self.lib.dll.hb_storage_num_global_dofs(self.h, level))

    def exchange_info(self, level):
        """(neighbouring ranks, doubles sent, doubles received) of one ghost exchange on `level`."""
        n = C.c_int(0)
        a, b = C.c_longlong(0), C.c_longlong(0)
        self.lib.check(self.lib.dll.hb_storage_exchange_info(self.h, level, C.byref(n), C.byref(a), C.byref(b)))
        return n.value, a.value, b.value

    def micro_coordinates(self, level):
        n = self.num_local_cells * self.cell_array_size(level)
        xyz = np.zeros((n, 3))
        self.lib.check(self.lib.dll.hb_storage_micro_coordinates(self.h, level, as_dp(xyz)))
        return xyz


class P1Function:
    def __init__(self, name, storage, min_level, max_level):
        self.name, self.storage, self.lib = name, storage, storage.lib
        self.min_level, self.max_level = min_level, max_level
        self.h = self.lib.checkp(self.lib.dll.hb_p1function_create(storage.h, min_level, max_level))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_p1function_destroy(self.h)
            self.h = None

    def _n(self, level):
        return self.storage.num_local_cells * self.storage.cell_array_size(level)

    def setBoundaryCondition(self, types):
        t = i32(types)
        self.lib.check(self.lib.dll.hb_p1function_set_boundary_condition(self.h, as_ip(t), len(t)))

    def interpolate(self, expr, level, flag=All):
        if callable(expr):
            xyz = self.storage.micro_coordinates(level)
            vals = f64(expr(xyz[:, 0], xyz[:, 1], xyz[:, 2]) * np.ones(len(xyz)))
            self.upload(vals, level, flag)
        else:
            self.lib.check(self.lib.dll.hb_p1function_interpolate_constant(self.h, float(expr), level, flag))

    def setToZero(self, level):
        self.interpolate(0.0, level, All)

    def upload(self, values, level, flag=All):
        v = f64(values).ravel()
        assert v.size == self._n(level), (v.size, self._n(level))
        self.lib.check(self.lib.dll.hb_p1function_upload(self.h, level, as_dp(v), flag))

    def download(self, level):
        out = np.zeros(self._n(level))
        self.lib.check(self.lib.dll.hb_p1function_download(self.h, level, as_dp(out)))
        return out

    def communicate(self, level):
        """explicit ghost exchange (operators do it themselves when needed)"""
        self.lib.check(self.lib.dll.hb_p1function_communicate(self.h, level))

    def device_ptr(self, level):
        return self.lib.dll.hb_p1function_device_ptr(self.h, level)

    def _handles(self, fns):
        arr = (C.c_void_p * len(fns))(*[f.h for f in fns])
        return arr

    def assign(self, scalars, functions, level, flag=All):
        s = f64(scalars)
        self.lib.check(self.lib.dll.hb_p1function_assign(self.h, len(functions), as_dp(s), self._handles(functions), level, flag))

    def add(self, scalars, *args):
        """add( scalars, functions, level, flag=All )  or  add( scalar, level, flag=All )"""
        if np.isscalar(scalars):
            level = args[0]
            flag = args[1] if len(args) > 1 else All
            self.lib.check(self.lib.dll.hb_p1function_add_scalar(self.h, float(scalars), level, flag))
            return
        functions, level = args[0], args[1]
        flag = args[2] if len(args) > 2 else All
        s = f64(scalars)
        self.lib.check(self.lib.dll.hb_p1function_add(self.h, len(functions), as_dp(s), self._handles(functions), level, flag))

    def multElementwise(self, functions, level, flag=All):
        self.lib.check(self.lib.dll.hb_p1function_mult_elementwise(self.h, len(functions), self._handles(functions), level, flag))

    def dotGlobal(self, other, level, flag=All):
        out = C.c_double()
        self.lib.check(self.lib.dll.hb_p1function_dot(self.h, other.h, level, flag, C.byref(out)))
        return out.value

    def sumGlobal(self, level, flag=All):
        out = C.c_double()
        self.lib.check(self.lib.dll.hb_p1function_sum(self.h, level, flag, C.byref(out)))
        return out.value

    def getMaxDoFMagnitude(self, level, flag=All):
        out = C.c_double()
        self.lib.check(self.lib.dll.hb_p1function_max_magnitude(self.h, level, flag, C.byref(out)))
        return out.value


class P1ConstantLaplaceOperator:
    def __init__(self, storage, min_level, max_level):
        self.storage, self.lib = storage, storage.lib
        self.min_level, self.max_level = min_level, max_level
        self.h = self.lib.checkp(self.lib.dll.hb_p1_constant_laplace_create(storage.h, min_level, max_level))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_operator_destroy(self.h)
            self.h = None

    def apply(self, src, dst, level, flag, update_type=Replace):
        self.lib.check(self.lib.dll.hb_operator_apply(self.h, src.h, dst.h, level, flag, update_type))

    def smooth_jac(self, dst, rhs, src, relax, level, flag):
        self.lib.check(self.lib.dll.hb_operator_smooth_jac(self.h, dst.h, rhs.h, src.h, relax, level, flag))

    def smooth_sor(self, dst, rhs, relax, level, flag, backwards=False, ordering=0):
        self.lib.check(self.lib.dll.hb_operator_smooth_sor_ordered(self.h, dst.h, rhs.h, relax, level, flag, int(backwards), ordering))

    def smooth_gs(self, dst, rhs, level, flag):
        self.smooth_sor(dst, rhs, 1.0, level, flag)

    def residual(self, x, b, r, level, flag):
        self.lib.check(self.lib.dll.hb_operator_residual(self.h, x.h, b.h, r.h, level, flag))

    def cheb_first(self, x, b, dst, level, flag):
        self.lib.check(self.lib.dll.hb_operator_cheb_first(self.h, x.h, b.h, dst.h, level, flag))

    def cheb_step(self, src, dst, x, coeff, level, flag):
        self.lib.check(self.lib.dll.hb_operator_cheb_step(self.h, src.h, dst.h, x.h, coeff, level, flag))

    def diagonal(self, f, level, inverse=False):
        self.lib.check(self.lib.dll.hb_operator_diagonal(self.h, f.h, level, int(inverse)))

    def cell_stencil(self, level, local_cell, mask=0):
        w = np.zeros(15)
        self.lib.check(self.lib.dll.hb_operator_cell_stencil(self.h, level, local_cell, mask, as_dp(w)))
        return w


class P1toP1LinearRestriction:
    def restrict(self, function, source_level, flag):
        function.lib.check(function.lib.dll.hb_p1_restrict(function.h, source_level, flag))


class P1toP1LinearProlongation:
    def prolongate(self, function, source_level, flag):
        function.lib.check(function.lib.dll.hb_p1_prolongate(function.h, source_level, flag))

    def prolongateAndAdd(self, function, source_level, flag):
        function.lib.check(function.lib.dll.hb_p1_prolongate_and_add(function.h, source_level, flag))


class Solver:
    def __init__(self, lib, handle, keep=()):
        self.lib, self.h, self._keep = lib, lib.checkp(handle), keep

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_solver_destroy(self.h)
            self.h = None

    def solve(self, A, x, b, level):
        self.lib.check(self.lib.dll.hb_solver_solve(self.h, A.h, x.h, b.h, level))


class CGSolver(Solver):
    def __init__(self, storage, min_level, max_level, max_iter=-1, rel_tol=1e-16, abs_tol=1e-16):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_cg(storage.h, min_level, max_level, max_iter, rel_tol, abs_tol), (storage,))

    def iterations(self):
        return self.lib.dll.hb_solver_iterations(self.h)


class WeightedJacobiSmoother(Solver):
    def __init__(self, storage, min_level, max_level, relax):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_weighted_jacobi(storage.h, min_level, max_level, relax), (storage,))


class SORSmoother(Solver):
    def __init__(self, storage, relax):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_sor(relax), (storage,))


class GaussSeidelSmoother(Solver):
    def __init__(self, storage):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_gauss_seidel(), (storage,))


class SymmetricGaussSeidelSmoother(Solver):
    def __init__(self, storage):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_symmetric_gauss_seidel(), (storage,))


class ChebyshevSmoother(Solver):
    def __init__(self, storage, min_level, max_level):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_chebyshev(storage.h, min_level, max_level), (storage,))

    def setupCoefficients(self, order, spectral_radius, upper_factor=1.2, lower_factor=0.3):
        self.setupCoefficientsInternal(order, lower_factor * spectral_radius, upper_factor * spectral_radius)

    def setupCoefficientsInternal(self, order, lower, upper):
        self.lib.check(self.lib.dll.hb_chebyshev_setup_coefficients(self.h, order, lower, upper))

    def coefficients(self):
        c = np.zeros(64)
        n = C.c_int()
        self.lib.check(self.lib.dll.hb_chebyshev_coefficients(self.h, as_dp(c), C.byref(n)))
        return c[: n.value].copy()


def estimateRadius(A, level, max_iter, x, tmp):
    """chebyshev::estimateRadius: power iteration on D^-1 A (x is the start vector, overwritten)."""
    out = C.c_double()
    A.lib.check(A.lib.dll.hb_estimate_spectral_radius(A.h, x.h, tmp.h, max_iter, level, C.byref(out)))
    return out.value


class DirectLevel0Solver(Solver):
    """dense Cholesky of the level-0 operator on the free macro vertices (unpartitioned storage); the reference's PETScLUSolver role"""

    def __init__(self, storage):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_direct_level0(storage.h), (storage,))


class ReplicatedCoarseCycle(Solver):
    """coarse-grid solver of a GeometricMultigridSolver on levels coarse_level .. L of a partitioned storage: the levels 0 .. coarse_level of
    the cycle (V(pre, post), Gauss-Seidel or weighted Jacobi, CG on level 0 -- cg_max_iter = 0: direct solve) on a replica of the whole
    mesh, redundantly on every rank"""

    def __init__(self, storage, coarse_level, smoother="gs", omega=0.66, pre=3, post=3, cg_max_iter=50, cg_rel_tol=1e-16, cg_abs_tol=1e-16):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_replicated_coarse_cycle(
            storage.h, coarse_level, 1 if smoother == "jacobi" else 0, omega, pre, post, cg_max_iter, cg_rel_tol, cg_abs_tol), (storage,))


class GeometricMultigridSolver(Solver):
    def __init__(self, storage, smoother, coarse_solver, min_level, max_level, pre=3, post=3, increment=0, cycle=VCYCLE):
        super().__init__(storage.lib, storage.lib.dll.hb_solver_create_gmg(
            storage.h, smoother.h, coarse_solver.h, min_level, max_level, pre, post, increment, cycle),
            (storage, smoother, coarse_solver))


# ---- P2 function space, blending, fused elementwise operators (BASELINE configs 3 / 4) ------------------------------------------------

def setIcosahedralShellMap(storage):
    """IcosahedralShellMap::setMap( setupStorage ) (geometry/IcosahedralShellMap.hpp:491-510)"""
    storage.lib.check(storage.lib.dll.hb_storage_set_icosahedral_shell_map(storage.h))


def shellMapParameters(storage, local_cell):
    out = np.zeros(14)
    storage.lib.check(storage.lib.dll.hb_storage_shell_map_parameters(storage.h, local_cell, as_dp(out)))
    return out


def evalMap(storage, level, local_cell, x):
    """GeometryMap::evalF / evalDF of a macro-cell at points x[n,3]: (F [n,3], DF [n,3,3])"""
    x = f64(x).reshape(-1, 3)
    out = np.zeros((len(x), 12))
    storage.lib.check(storage.lib.dll.hb_storage_map_eval(storage.h, level, local_cell, as_dp(x), len(x), as_dp(out)))
    return out[:, :3].copy(), out[:, 3:].reshape(-1, 3, 3).copy()


class _VertexPart(P1Function):
    """P2Function::getVertexDoFFunction(): a view on the P2 function's vertex DoFs (owned by the P2 function)"""

    def __init__(self, p2, handle):
        self.name, self.storage, self.lib = p2.name + "_v", p2.storage, p2.lib
        self.min_level, self.max_level = p2.min_level, p2.max_level
        self.h = handle
        self._owner = p2

    def __del__(self):
        self.h = None


class P2Function:
    def __init__(self, name, storage, min_level, max_level):
        self.name, self.storage, self.lib = name, storage, storage.lib
        self.min_level, self.max_level = min_level, max_level
        self.h = self.lib.checkp(self.lib.dll.hb_p2function_create(storage.h, min_level, max_level))
        self.v = _VertexPart(self, self.lib.dll.hb_p2function_vertex_part(self.h))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_p2function_destroy(self.h)
            self.h = None

    def getVertexDoFFunction(self):
        return self.v

    def edge_size(self, level):
        return self.storage.num_local_cells * int(self.lib.dll.hb_p2_edge_cell_size(level))

    def upload(self, vertex_values, edge_values, level):
        self.v.upload(vertex_values, level, All)
        e = f64(edge_values).ravel()
        assert e.size == self.edge_size(level), (e.size, self.edge_size(level))
        self.lib.check(self.lib.dll.hb_p2function_upload_edges(self.h, level, as_dp(e)))

    def download(self, level):
        e = np.zeros(self.edge_size(level))
        self.lib.check(self.lib.dll.hb_p2function_download_edges(self.h, level, as_dp(e)))
        return self.v.download(level), e

    def interpolate(self, value, level, flag=All):
        self.lib.check(self.lib.dll.hb_p2function_interpolate_constant(self.h, float(value), level, flag))

    def _lin(self, fn, scalars, functions, level, flag):
        s = f64(scalars)
        hs = (C.c_void_p * len(functions))(*[f.h for f in functions])
        self.lib.check(fn(self.h, len(functions), as_dp(s), hs, level, flag))

    def assign(self, scalars, functions, level, flag=All):
        self._lin(self.lib.dll.hb_p2function_assign, scalars, functions, level, flag)

    def add(self, scalars, functions, level, flag=All):
        self._lin(self.lib.dll.hb_p2function_add, scalars, functions, level, flag)

    def dotGlobal(self, other, level, flag=All):
        out = C.c_double(0)
        self.lib.check(self.lib.dll.hb_p2function_dot(self.h, other.h, level, flag, C.byref(out)))
        return out.value


class P2ElementwiseOperator:
    """P2ElementwiseOperator< Form > (elementwiseoperators/P2ElementwiseOperator.hpp); k: P2Function, callable k(x, y, z) on
    physical points (sampled at the quadrature points, the in-tree form's std::function), or None (k = 1)"""

    def __init__(self, storage, min_level, max_level, handle, k):
        self.storage, self.lib = storage, storage.lib
        self.min_level, self.max_level = min_level, max_level
        self.h = self.lib.checkp(handle)
        self._k = k
        if callable(k):
            for L in range(min_level, max_level + 1):
                n = int(self.lib.dll.hb_p2operator_num_quadrature_points(self.h, L))
                xyz = np.zeros((n, 3))
                self.lib.check(self.lib.dll.hb_p2operator_quadrature_points(self.h, L, as_dp(xyz)))
                vals = f64(k(xyz[:, 0], xyz[:, 1], xyz[:, 2]) * np.ones(n))
                self.lib.check(self.lib.dll.hb_p2operator_set_sampled_coefficient(self.h, L, as_dp(vals)))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_p2operator_destroy(self.h)
            self.h = None

    def apply(self, src, dst, level, flag, update=Replace):
        self.lib.check(self.lib.dll.hb_p2operator_apply(self.h, src.h, dst.h, level, flag, update))


class P2ElementwiseBlendingDivKGradOperator(P2ElementwiseOperator):
    def __init__(self, storage, min_level, max_level, k=None):
        kh = k.h if isinstance(k, P2Function) else None
        super().__init__(storage, min_level, max_level, storage.lib.dll.hb_p2_divkgrad_create(storage.h, min_level, max_level, kh), k)


class P2ElementwiseBlendingEpsilonBlock(P2ElementwiseOperator):
    """block (i, j) of P2ElementwiseBlendingEpsilonOperator (elementwiseoperators/P2ElementwiseEpsilonOperator.hpp:112-170)"""

    def __init__(self, storage, min_level, max_level, i, j, mu=None):
        kh = mu.h if isinstance(mu, P2Function) else None
        super().__init__(storage, min_level, max_level, storage.lib.dll.hb_p2_epsilon_create(storage.h, min_level, max_level, i, j, kh), mu)


class P2toP2QuadraticRestriction:
    def restrict(self, f, source_level, flag):
        f.lib.check(f.lib.dll.hb_p2_restrict(f.h, source_level, flag))


class P2toP2QuadraticProlongation:
    def prolongate(self, f, source_level, flag):
        f.lib.check(f.lib.dll.hb_p2_prolongate(f.h, source_level, flag))

    def prolongateAndAdd(self, f, source_level, flag):
        f.lib.check(f.lib.dll.hb_p2_prolongate_and_add(f.h, source_level, flag))


class P2ToP1ElementwiseBlendingDivOperator(P2ElementwiseOperator):
    """P2ToP1ElementwiseBlendingDiv{x,y,z}Operator (elementwiseoperators/P2ToP1ElementwiseOperator.hpp:264-266)"""

    def __init__(self, storage, min_level, max_level, component):
        super().__init__(storage, min_level, max_level, storage.lib.dll.hb_p2_div_create(storage.h, min_level, max_level, component), None)

    def apply(self, src, dst, level, flag, update=Replace):
        self.lib.check(self.lib.dll.hb_p2operator_apply_p2_to_p1(self.h, src.h, dst.h, level, flag, update))


class P1ToP2ElementwiseBlendingDivTOperator(P2ElementwiseOperator):
    """P1ToP2ElementwiseBlendingDivT{x,y,z}Operator (elementwiseoperators/P1ToP2ElementwiseOperator.hpp:275-277)"""

    def __init__(self, storage, min_level, max_level, component):
        super().__init__(storage, min_level, max_level, storage.lib.dll.hb_p2_divt_create(storage.h, min_level, max_level, component), None)

    def apply(self, src, dst, level, flag, update=Replace):
        self.lib.check(self.lib.dll.hb_p2operator_apply_p1_to_p2(self.h, src.h, dst.h, level, flag, update))


class P2ElementwiseBlendingEpsilonOperator(P2ElementwiseOperator):
    """P2ElementwiseBlendingEpsilonOperator (elementwiseoperators/P2ElementwiseEpsilonOperator.hpp:112-170) on a P2 vector function
    (a list of three P2Functions), all nine blocks in one fused kernel; mu: P2Function, callable or None"""

    def __init__(self, storage, min_level, max_level, mu=None):
        kh = mu.h if isinstance(mu, P2Function) else None
        super().__init__(storage, min_level, max_level, storage.lib.dll.hb_p2_epsilon_vector_create(storage.h, min_level, max_level, kh), mu)

    def apply(self, src, dst, level, flag, update=Replace):
        hs = (C.c_void_p * 3)(*[f.h for f in src])
        hd = (C.c_void_p * 3)(*[f.h for f in dst])
        self.lib.check(self.lib.dll.hb_p2operator_apply_vector(self.h, hs, hd, level, flag, update))


class P2P1TaylorHoodFunction:
    """P2P1TaylorHoodFunction (composites/P2P1TaylorHoodFunction.hpp): uvw = three P2Functions, p = a P1Function"""

    def __init__(self, name, storage, min_level, max_level):
        self.uvw = [P2Function("%s_%s" % (name, c), storage, min_level, max_level) for c in "uvw"]
        self.p = P1Function(name + "_p", storage, min_level, max_level)


class P2P1ElementwiseBlendingEpsilonStokesOperator:
    """The Taylor-Hood Stokes operator with the epsilon (full viscous) velocity block: A u + B^T p ; B u -- the composite of
    P2P1ElementwiseBlendingStokesOperator::apply (elementwiseoperators/P2P1ElementwiseBlendingStokesOperator.hpp:91-100) with
    P2ElementwiseBlendingEpsilonOperator as A (BASELINE config 4; src/hyteg_operators_composites/stokes/P2P1StokesEpsilonOperator.hpp)"""

    def __init__(self, storage, min_level, max_level, mu=None, fused=True):
        self.A = P2ElementwiseBlendingEpsilonOperator(storage, min_level, max_level, mu)
        self.div = [P2ToP1ElementwiseBlendingDivOperator(storage, min_level, max_level, c) for c in range(3)]
        self.divT = [P1ToP2ElementwiseBlendingDivTOperator(storage, min_level, max_level, c) for c in range(3)]
        # one kernel for A u + B^T p and B u: the geometry of a quadrature point is evaluated once instead of seven times
        kh = mu.h if isinstance(mu, P2Function) else None
        self.fused = P2ElementwiseOperator(storage, min_level, max_level, storage.lib.dll.hb_p2_stokes_create(storage.h, min_level, max_level, kh), mu) if fused else None

    def apply(self, src, dst, level, flag, update=Replace):
        assert update == Replace
        if self.fused is not None:
            hs = (C.c_void_p * 3)(*[f.h for f in src.uvw])
            hd = (C.c_void_p * 3)(*[f.h for f in dst.uvw])
            self.fused.lib.check(self.fused.lib.dll.hb_p2operator_apply_stokes(self.fused.h, hs, src.p.h, hd, dst.p.h, level, flag))
            return
        self.A.apply(src.uvw, dst.uvw, level, flag, Replace)
        for c in range(3):
            self.divT[c].apply(src.p, dst.uvw[c], level, flag, Add)
            self.div[c].apply(src.uvw[c], dst.p, level, flag, Replace if c == 0 else Add)


class P2ChebyshevGMG:
    """GeometricMultigridSolver< P2 elementwise operator > with ChebyshevSmoother, quadratic transfer and CG on the coarsest level"""

    def __init__(self, storage, A, min_level, max_level, order, ev_guess, power_iterations=20, pre=2, post=2, coarse_max_iter=1000, coarse_tol=1e-14):
        self.lib, self._keep = storage.lib, (storage, A, ev_guess)
        self.h = self.lib.checkp(self.lib.dll.hb_p2_gmg_chebyshev_create(storage.h, A.h, min_level, max_level, order, ev_guess.h, power_iterations,
                                                                         pre, post, coarse_max_iter, coarse_tol))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.dll.hb_p2solver_destroy(self.h)
            self.h = None

    @property
    def spectral_radius(self):
        return self.lib.dll.hb_p2solver_spectral_radius(self.h)

    def solve(self, x, b, level):
        self.lib.check(self.lib.dll.hb_p2solver_solve(self.h, x.h, b.h, level, 0))

    def smooth(self, x, b, level):
        self.lib.check(self.lib.dll.hb_p2solver_solve(self.h, x.h, b.h, level, 1))
