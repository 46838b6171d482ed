// Forced include (g++ -include) that compiles an UNMODIFIED reference driver or test with hyteg::P1ConstantLaplaceOperator replaced by the
// device bridge: the reference's own header is read first (its #pragma once makes the driver's later #include a no-op), then every later
// mention of the class name -- the operator object, the GaussSeidelSmoother / CGSolver / GeometricMultigridSolver template arguments -- is
// rewritten to the bridge class.  Used by oracle/build_ref.py to build tests/hyteg/convergence/P1GMG3DConvergenceTest.cpp of the
// reference against libhyteg_b200.so without touching its source.
#pragma once
#include "constant_stencil_operator/P1ConstantOperator.hpp"

#include "P1ConstantLaplaceOperatorB200.hpp"

#define P1ConstantLaplaceOperator P1ConstantLaplaceOperatorB200
