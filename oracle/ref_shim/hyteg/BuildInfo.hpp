// Hand-written stand-in for src/hyteg/BuildInfo.in.hpp (configure-time header).
#pragma once
#include <string>
namespace hyteg::buildinfo {
inline std::string buildType() { return "Release"; }
inline std::string compilerFlags() { return "-O3"; }
inline std::string compilerInfo() { return "g++"; }
inline std::string systemEndianess() { return "LITTLE_ENDIAN"; }
inline std::string mpiVersion() { return "no MPI library info; sequential build!"; }
inline std::string fpType() { return "double"; }
inline void printBuildInfo() {}
}
