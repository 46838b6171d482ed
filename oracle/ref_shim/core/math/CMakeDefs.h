// Hand-written stand-in for walberla/src/core/math/CMakeDefs.in.h (configure-time header).
#pragma once
