// Hand-written stand-in for walberla/src/core/waLBerlaBuildInfo.in.cpp (configure-time source).
#include "core/waLBerlaBuildInfo.h"
namespace walberla { namespace core { namespace buildinfo {
const char* gitSHA1() { return "b30bbaf4"; }
const char* buildType() { return "Release"; }
const char* compilerFlags() { return "-O3"; }
const char* buildMachine() { return "oracle/build_ref.py"; }
const char* sourceDir() { return "/root/reference/walberla"; }
const char* binaryDir() { return "oracle/_ref"; }
}}}
