// CUDA / host-emulation portability shim.
//
// The product library is compiled by nvcc for sm_100a and nothing else.  For CPU-only CI (this
// repository is developed on a box without a GPU) the SAME kernel source can be compiled by g++
// with -DHB_EMUL into a TEST-ONLY library (tests/emul/): kernels then run as sequential loops over
// blocks and threads.  The emulation library is never loaded by the hyteg_b200 package -- it exists
// so that the host logic (topology tables, orientation maps, solver drivers) can be unit-tested
// without a device, in the same spirit as HyTeG's non-MPI stub build (reference CMakeLists.txt:51-55).
#pragma once

#ifdef HB_EMUL

#include <cmath>
#include <cstdlib>
#include <cstring>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__( ... )
#define HB_CONSTANT static const
#define HB_WARP 1

struct hbDim3
{
   unsigned x, y, z;
   hbDim3( unsigned a = 1, unsigned b = 1, unsigned c = 1 )
   : x( a )
   , y( b )
   , z( c )
   {}
};
typedef hbDim3 dim3;
extern dim3    threadIdx, blockIdx, blockDim, gridDim;

template < typename T >
inline T __ldg( const T* p )
{
   return *p;
}

typedef int   cudaError_t;
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
#define cudaSuccess 0
enum cudaMemcpyKind
{
   cudaMemcpyHostToDevice,
   cudaMemcpyDeviceToHost,
   cudaMemcpyDeviceToDevice
};
inline cudaError_t cudaMalloc( void** p, size_t n )
{
   *p = std::malloc( n ? n : 1 );
   return *p ? 0 : 2;
}
inline cudaError_t cudaFree( void* p )
{
   std::free( p );
   return 0;
}
inline cudaError_t cudaMallocHost( void** p, size_t n )
{
   return cudaMalloc( p, n );
}
inline cudaError_t cudaFreeHost( void* p )
{
   return cudaFree( p );
}
inline cudaError_t cudaMemcpyAsync( void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t )
{
   std::memmove( d, s, n );
   return 0;
}
inline cudaError_t cudaMemcpy( void* d, const void* s, size_t n, cudaMemcpyKind )
{
   std::memmove( d, s, n );
   return 0;
}
inline cudaError_t cudaMemset( void* d, int v, size_t n )
{
   std::memset( d, v, n );
   return 0;
}
inline cudaError_t cudaMemsetAsync( void* d, int v, size_t n, cudaStream_t )
{
   std::memset( d, v, n );
   return 0;
}
inline cudaError_t cudaStreamSynchronize( cudaStream_t )
{
   return 0;
}
inline cudaError_t cudaDeviceSynchronize()
{
   return 0;
}
inline cudaError_t cudaGetLastError()
{
   return 0;
}
inline cudaError_t cudaSetDevice( int )
{
   return 0;
}
inline const char* cudaGetErrorString( cudaError_t )
{
   return "emulation";
}

// sequential launch: one "thread" per block (kernels are written blockDim-agnostic)
#define HB_LAUNCH( kernel, grid, block, stream, ... )                      \
   do                                                                      \
   {                                                                       \
      gridDim  = dim3( ( grid ) );                                         \
      blockDim = dim3( ( block ) );                                        \
      for ( unsigned hb_b = 0; hb_b < gridDim.x; hb_b++ )                  \
         for ( unsigned hb_t = 0; hb_t < blockDim.x; hb_t++ )              \
         {                                                                 \
            blockIdx  = dim3( hb_b );                                      \
            threadIdx = dim3( hb_t );                                      \
            kernel( __VA_ARGS__ );                                         \
         }                                                                 \
   } while ( 0 )

#else // real CUDA

#include <cuda_runtime.h>
#define HB_CONSTANT __constant__ const
#define HB_WARP 32
#define HB_LAUNCH( kernel, grid, block, stream, ... ) kernel<<< ( grid ), ( block ), 0, ( stream ) >>>( __VA_ARGS__ )

#endif
