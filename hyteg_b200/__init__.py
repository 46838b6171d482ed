"""hyteg_b200: B200-native matrix-free P1 geometric multigrid behind HyTeG's operator API.

Product path = CUDA library (hyteg_b200/libhyteg_b200.so) through the C-ABI in include/hyteg_b200.h.
"""
from .capi import (Add, All, Boundary, DirichletBoundary, FreeslipBoundary, HytegB200Error, Inner,
                   NeumannBoundary, Replace, VCYCLE, WCYCLE, get_lib)
from .api import (CGSolver, ChebyshevSmoother, GaussSeidelSmoother, GeometricMultigridSolver,
                  P1ConstantLaplaceOperator, P1Function, P1toP1LinearProlongation, P1toP1LinearRestriction,
                  PrimitiveStorage, SMOOTH_FLAG, SORSmoother, SymmetricGaussSeidelSmoother,
                  WeightedJacobiSmoother, estimateRadius)
from . import meshes
