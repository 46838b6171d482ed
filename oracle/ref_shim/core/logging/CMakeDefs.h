// Hand-written stand-in for walberla/src/core/logging/CMakeDefs.in.h (configure-time header).
#pragma once
#define WALBERLA_LOGLEVEL_INFO
#define WALBERLA_LOGLEVEL INFO
#define WALBERLA_LOGLEVEL_STRING "INFO"
