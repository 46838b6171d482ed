// Hand-written stand-in for the header waLBerla's CMake configure step would generate
// (walberla/src/waLBerlaDefinitions.in.h).  Serial, double precision, no optional packages.
#pragma once
#define WALBERLA_DOUBLE_ACCURACY
#define WALBERLA_CXX_COMPILER_IS_GNU
#define WALBERLA_BUILD_WITH_BOOST_THREAD 0
