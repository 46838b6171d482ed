// Hand-written stand-in for src/hyteg/Git.in.hpp (configure-time header).
#pragma once
#include <string>
namespace hyteg::buildinfo {
inline std::string gitSHA1() { return "a7ecfdf6"; }
inline std::string gitBranch() { return "reference"; }
inline std::string gitDiff() { return ""; }
inline void printGitInfo() {}
}
