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
// mesh files are always handed to the oracle driver by explicit path; the macro only has to expand to a string
#ifndef HYTEG_MESH_DIR
#define HYTEG_MESH_DIR "/root/reference/data/meshes"
#endif
