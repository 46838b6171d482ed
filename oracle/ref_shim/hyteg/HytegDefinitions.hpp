// Hand-written stand-in for the header HyTeG's CMake configure step would generate
// (src/hyteg/HytegDefinitions.in.hpp).  Serial build, generated kernels on, no optional packages.
#pragma once
#define HYTEG_USE_GENERATED_KERNELS
namespace hyteg {
namespace globalDefines {
constexpr bool useGeneratedKernels = true;
}
} // namespace hyteg
#define HYTEG_ARCH_ENDIANESS LITTLE_ENDIAN
