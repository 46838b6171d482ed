// TMA-streamed interior stencil kernel (sm_100a): the hot loop of apply / residual / Jacobi / Chebyshev.
//
// Replaces apply_3D_macrocell_vertexdof_to_vertexdof_{replace,add} (reference
// src/constant_stencil_operator/P1generatedKernels/apply_3D_macrocell_vertexdof_to_vertexdof_replace.cpp:52-77)
// and the BLAS-1 passes the reference chains behind it (P1Operator.hpp:429-447, ChebyshevSmoother.hpp:179-222).
//
// Structure:
//   * persistent CTAs (2 per SM) pull work items {macro-cell, 8 z-slices, 64 x-columns, <=30 rows} from an atomic counter;
//   * one producer warp streams the item row by row: for every input row y and each of the 10 slices z0-1..z0+8 one
//     cp.async.bulk (1D TMA, UBLKCP) copies the 68-double row segment into a ring of shared-memory stages, completion is
//     signalled on an mbarrier (complete_tx); extra operands of the fused epilogue (rhs / dst / x) ride in the same stage;
//   * 8 consumer warps = 4 slice pairs x 2 half-tiles.  Every thread owns one x column and TWO z slices, marches along y
//     and keeps the three live rows of its four input slices in registers: 12 conflict-free LDS.64 + 30 DFMA per 2 outputs;
//   * results leave through coalesced STG.64.
// Every source element is fetched 10/8 (z halo) x 68/64 (x halo + 16 B alignment) x 32/30 (y restart) = 1.42x from L2
// and, with neighbouring items running concurrently, once from HBM.
#include "stream_kernel.h"

#ifndef HB_EMUL

#include <cstdint>

namespace {

constexpr int NCONS   = 8;   // consumer warps = 2 warpgroups
constexpr int THREADS = 384; // + 1 warpgroup: producer warp and three idle warps that hand their registers back
constexpr int REGS_CONSUMER = 104; // 8 x 104 + 4 x 24 <= 12 x 80: setmaxnreg moves registers inside ONE CTA only
constexpr int REGS_PRODUCER = 24;

// geometry of a work item for XW warps along x: ZB output slices x XT output columns
template < int XW >
struct Geo
{
   static constexpr int ZB       = 2 * ( NCONS / XW );
   static constexpr int XT       = 32 * XW;
   static constexpr int ROWLEN   = XT + 4; // staged doubles per row segment (2 halo + 2 alignment slack)
   static constexpr int ROWBYTES = ROWLEN * 8;
   static constexpr int NSRC     = ZB + 2;
};

__device__ __forceinline__ uint32_t smem_u32( const void* p )
{
   return (uint32_t) __cvta_generic_to_shared( p );
}
__device__ __forceinline__ void mbar_init( uint64_t* bar, int count )
{
   asm volatile( "mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"( smem_u32( bar ) ), "r"( count ) : "memory" );
}
__device__ __forceinline__ void mbar_arrive( uint64_t* bar )
{
   asm volatile( "{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"( smem_u32( bar ) ) : "memory" );
}
__device__ __forceinline__ void mbar_arrive_expect_tx( uint64_t* bar, uint32_t bytes )
{
   asm volatile( "{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"( smem_u32( bar ) ), "r"( bytes ) : "memory" );
}
__device__ __forceinline__ void mbar_wait( uint64_t* bar, uint32_t parity )
{
   asm volatile( "{\n"
                 ".reg .pred p;\n"
                 "WAIT_LOOP:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra WAIT_DONE;\n"
                 "bra WAIT_LOOP;\n"
                 "WAIT_DONE:\n"
                 "}" ::"r"( smem_u32( bar ) ),
                 "r"( parity )
                 : "memory" );
}
__device__ __forceinline__ void tma_load_1d( void* smemDst, const void* gmemSrc, uint32_t bytes, uint64_t* bar )
{
   asm volatile( "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"( smem_u32( smemDst ) ),
                 "l"( gmemSrc ),
                 "r"( bytes ),
                 "r"( smem_u32( bar ) )
                 : "memory" );
}

__device__ __forceinline__ int tet( int k )
{
   return ( k * ( k + 1 ) * ( k + 2 ) ) / 6;
}
// index of (x=0, y, z) inside a macro-cell array
__device__ __forceinline__ int row_base( int N, int y, int z )
{
   const int wz = N + 1 - z;
   return tet( N + 1 ) - tet( wz ) + y * ( wz + 1 ) - ( y * ( y + 1 ) ) / 2;
}

struct StageHeader
{
   int32_t cell; // -1: no more work
   int32_t z0, x0;
   int32_t yIn;   // first input row of this chunk
   int32_t yOut0; // first output row of the item
   int32_t yOut1; // last output row of the item
   int32_t pad[2];
};

template < int MODE >
struct ModeTraits
{
   static constexpr int  nExtra   = ( MODE == HB_APPLY_REPLACE ) ? 0 : 1;
   static constexpr bool extraRhs = ( MODE == HB_RESIDUAL || MODE == HB_JACOBI || MODE == HB_CHEB_FIRST );
};

} // namespace

namespace {

template < int MODE >
__device__ __forceinline__ void hb_epilogue_stream( const HbStencilArgs& a, size_t gi, double sNb, double diag, double invd, double centre, double e )
{
   // same operation order as hb_epilogue (hb_device.h): results are bit-identical to the generic kernel
   const double Ax = sNb + diag * centre;
   if ( MODE == HB_APPLY_REPLACE )
      a.dst[gi] = Ax;
   else if ( MODE == HB_APPLY_ADD )
      a.dst[gi] = e + Ax;
   else if ( MODE == HB_RESIDUAL )
      a.dst[gi] = e - Ax;
   else if ( MODE == HB_JACOBI )
      a.dst[gi] = centre + a.omega * ( invd * ( e - Ax ) );
   else if ( MODE == HB_CHEB_FIRST )
      a.dst[gi] = invd * ( e - Ax );
   else if ( MODE == HB_CHEB_STEP )
   {
      a.dst[gi] = invd * Ax;
      a.aux[gi] = e + a.omega * centre;
   }
}

template < int MODE, int XW, int R, int STAGES >
struct Smem
{
   using G                         = Geo< XW >;
   static constexpr int NR         = G::NSRC + G::ZB * ModeTraits< MODE >::nExtra; // staged row segments per input row
   static constexpr int stageBytes = R * NR * G::ROWBYTES;
   static constexpr int dataOff    = 1024;
   static constexpr int total      = dataOff + STAGES * stageBytes;
};

template < int MODE, int XW, int R, int STAGES >
__global__ void __launch_bounds__( THREADS, 2 ) hb_stream_kernel( const HbLevelView lv, const HbStencilArgs a, const HbStreamArgs sa )
{
   using SM = Smem< MODE, XW, R, STAGES >;
   using G  = Geo< XW >;
   constexpr int NR = SM::NR, ROWLEN = G::ROWLEN, NSRC = G::NSRC;
   extern __shared__ __align__( 128 ) unsigned char smemRaw[];
   StageHeader* hdr   = reinterpret_cast< StageHeader* >( smemRaw );    // STAGES * 32 B
   uint64_t*    full  = reinterpret_cast< uint64_t* >( smemRaw + 512 ); // STAGES * 8 B
   uint64_t*    empty = reinterpret_cast< uint64_t* >( smemRaw + 768 ); // STAGES * 8 B
   double*      data  = reinterpret_cast< double* >( smemRaw + SM::dataOff );

   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   const int N    = lv.N;
   if ( threadIdx.x == 0 )
   {
      for ( int s = 0; s < STAGES; s++ )
      {
         mbar_init( &full[s], 4 ); // one arrival (with its byte count) per producer warp
         mbar_init( &empty[s], NCONS );
      }
      asm volatile( "fence.mbarrier_init.release.cluster;" ::: "memory" );
   }
   __syncthreads();

   int      stage = 0;
   uint32_t phase = 0;

   if ( warp >= NCONS )
   {
      // ---------------------------------------------------- producer warpgroup: gives its registers to the consumers
      asm volatile( "setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"( REGS_PRODUCER ) );
      // Four producer warps share the copies of a stage (one row segment per lane): a single warp issuing all of them is
      // latency-bound (ncu r1_c: consumers starved, 65 % of the samples in the full-barrier wait).
      static_assert( R * NR <= 128, "one copy per producer lane" );
      const int  pw    = warp - NCONS;
      const int  k     = pw + 4 * lane; // copy handled by this lane: interleaved so that every warp gets a quarter
      const int  j     = k / NR, q = k - j * NR;
      const bool isSrc = q < NSRC;
      const int  total = lv.nCells * sa.nItems;
      int*       itemSlot = reinterpret_cast< int* >( smemRaw + 1000 );
      const double* base  = isSrc ? sa.srcArena : sa.extraArena;
      while ( true )
      {
         if ( threadIdx.x == NCONS * 32 )
            *itemSlot = atomicAdd( sa.counter, 1 );
         asm volatile( "bar.sync 1, 128;" ::: "memory" );
         const int item = *itemSlot;
         asm volatile( "bar.sync 2, 128;" ::: "memory" ); // slot may be overwritten after everybody has read it
         if ( item >= total )
            break;
         const int cell = item / sa.nItems;
         if ( !( ( sa.selMask >> lv.cells[cell].region[0].bflag ) & 1 ) )
            continue;
         const HbItem    it       = sa.items[item - cell * sa.nItems];
         const int       nChunks  = ( it.nOut + 2 + R - 1 ) / R;
         const long long cellBase = (long long) cell * lv.cellSize;
         // this lane's row segment: slice s, columns xa..xb, rows yFirst + c*R
         const int s      = isSrc ? it.z0 - 1 + q : it.z0 + ( q - NSRC );
         const int xa     = isSrc ? it.x0 - 1 : it.x0;
         const int xbWant = isSrc ? it.x0 + G::XT : it.x0 + G::XT - 1;
         const int yFirst = ( isSrc ? it.ya - 1 : it.ya - 2 ) + j;
         for ( int c = 0; c < nChunks; c++ )
         {
            mbar_wait( &empty[stage], phase ^ 1u );
            double*   sdata = data + (size_t) stage * ( SM::stageBytes / 8 );
            const int y     = yFirst + c * R;
            const int last  = N - s - y; // last column of the row
            const int xb    = xbWant < last ? xbWant : last;
            long long g0    = 0;
            int       bytes = 0;
            if ( k < R * NR && y >= 0 && s <= N && xb >= xa )
            {
               const long long g  = cellBase + row_base( N, y, s ) + xa;
               const long long g1 = ( g + ( xb - xa ) + 2 ) & ~1LL; // one past the last element, rounded up to even
               g0                 = g & ~1LL;
               bytes              = (int) ( g1 - g0 ) * 8;
            }
            int sum = bytes;
#pragma unroll
            for ( int o = 16; o > 0; o >>= 1 )
               sum += __shfl_xor_sync( 0xffffffffu, sum, o );
            if ( lane == 0 )
            {
               if ( pw == 0 )
               {
                  StageHeader h;
                  h.cell     = cell;
                  h.z0       = it.z0;
                  h.x0       = it.x0;
                  h.yIn      = it.ya - 1 + c * R;
                  h.yOut0    = it.ya;
                  h.yOut1    = it.ya + it.nOut - 1;
                  hdr[stage] = h;
               }
               mbar_arrive_expect_tx( &full[stage], (uint32_t) sum );
            }
            __syncwarp();
            if ( bytes > 0 )
               tma_load_1d( sdata + k * ROWLEN, base + g0, (uint32_t) bytes, &full[stage] );
            if ( ++stage == STAGES )
            {
               stage = 0;
               phase ^= 1u;
            }
         }
      }
      if ( pw == 0 )
      {
         mbar_wait( &empty[stage], phase ^ 1u );
         if ( lane == 0 )
            hdr[stage].cell = -1;
      }
      if ( lane == 0 )
      {
         mbar_arrive( &full[stage] );
         if ( pw == 0 && atomicAdd( sa.counter + 1, 1 ) == (int) gridDim.x - 1 )
         {
            // last CTA to run out of work re-arms the counters for the next launch
            sa.counter[0] = 0;
            sa.counter[1] = 0;
         }
      }
      return;
   }

   // --------------------------------------------------------------------- consumer warps
   asm volatile( "setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"( REGS_CONSUMER ) );
   const int p  = warp / XW;                          // slice pair
   const int xl = ( ( warp % XW ) << 5 ) + lane;      // local column
   int       curCell = -1;
   double    w[HB_NDIR];
   double    invd = 0.0;
   // register window: slice q (0..3) rows A (y-2), B (y-1); suffix l/c/r = column x-1/x/x+1
   double q0Bc = 0, q0Br = 0;
   double q1Ac = 0, q1Ar = 0, q1Bl = 0, q1Bc = 0, q1Br = 0;
   double q2Ac = 0, q2Ar = 0, q2Bl = 0, q2Bc = 0, q2Br = 0;
   double q3Ac = 0, q3Ar = 0, q3Bl = 0, q3Bc = 0, q3Br = 0;
   long long cellBase = 0;
   int       zA = 0, x = 0, cpar = 0, gi[4] = { 0, 0, 0, 0 }, ob0 = 0, ob1 = 0, yMax0 = -1, yMax1 = -1, yMin = 0;

   while ( true )
   {
      mbar_wait( &full[stage], phase );
      const StageHeader h = hdr[stage];
      if ( h.cell < 0 )
         break;
      if ( h.yIn == h.yOut0 - 1 )
      {
         // first chunk of a work item: (re)load the weights and derive the per-thread index state once per item
         if ( h.cell != curCell )
         {
            curCell                       = h.cell;
            const double* __restrict__ Wc = a.W + (size_t) h.cell * HB_NMASK * HB_NDIR;
#pragma unroll
            for ( int d = 0; d < HB_NDIR; d++ )
               w[d] = __ldg( Wc + d );
            invd = 1.0 / w[D_C];
         }
         cellBase = (long long) h.cell * lv.cellSize;
         zA       = h.z0 + 2 * p;
         x        = h.x0 + xl;
         cpar     = (int) ( cellBase & 1 );
#pragma unroll
         for ( int q = 0; q < 4; q++ )
            gi[q] = row_base( N, h.yIn, zA - 1 + q ) + h.x0 - 1 + cpar; // parity of the staged row segments
         ob0 = row_base( N, h.yIn - 1, zA ); // output row bases (row yIn-1 of slices zA, zA+1)
         ob1 = row_base( N, h.yIn - 1, zA + 1 );
         // last valid output row of this thread's column in its two slices (-1: none)
         yMax0 = ( zA <= N - 3 ) ? min( h.yOut1, min( N - 2 - zA, N - 1 - zA - x ) ) : -1;
         yMax1 = ( zA + 1 <= N - 3 ) ? min( h.yOut1, min( N - 3 - zA, N - 2 - zA - x ) ) : -1;
         yMin  = h.yOut0;
      }
      const double* sdata = data + (size_t) stage * ( SM::stageBytes / 8 );

#pragma unroll
      for ( int j = 0; j < R; j++ )
      {
         const int     yr  = h.yIn + j; // new input row
         const int     yo  = yr - 1;    // output row
         const double* row = sdata + ( j * NR + 2 * p ) * ROWLEN + xl;
         const double* r0  = row + ( gi[0] & 1 );
         const double* r1  = row + ROWLEN + ( gi[1] & 1 );
         const double* r2  = row + 2 * ROWLEN + ( gi[2] & 1 );
         const double* r3  = row + 3 * ROWLEN + ( gi[3] & 1 );
         const double  n0l = r0[0], n0c = r0[1], n0r = r0[2];
         const double  n1l = r1[0], n1c = r1[1], n1r = r1[2];
         const double  n2l = r2[0], n2c = r2[1], n2r = r2[2];
         const double  n3l = r3[0], n3c = r3[1], n3r = r3[2];

         // output (zA, yo, x): slices q0 (below), q1 (own), q2 (above)
         double s0 = w[D_W] * q1Bl;
         s0 += w[D_E] * q1Br;
         s0 += w[D_N] * n1c;
         s0 += w[D_NW] * n1l;
         s0 += w[D_S] * q1Ac;
         s0 += w[D_SE] * q1Ar;
         s0 += w[D_BC] * q0Bc;
         s0 += w[D_BE] * q0Br;
         s0 += w[D_BN] * n0c;
         s0 += w[D_BNW] * n0l;
         s0 += w[D_TC] * q2Bc;
         s0 += w[D_TW] * q2Bl;
         s0 += w[D_TS] * q2Ac;
         s0 += w[D_TSE] * q2Ar;
         // output (zA+1, yo, x): slices q1 (below), q2 (own), q3 (above)
         double s1 = w[D_W] * q2Bl;
         s1 += w[D_E] * q2Br;
         s1 += w[D_N] * n2c;
         s1 += w[D_NW] * n2l;
         s1 += w[D_S] * q2Ac;
         s1 += w[D_SE] * q2Ar;
         s1 += w[D_BC] * q1Bc;
         s1 += w[D_BE] * q1Br;
         s1 += w[D_BN] * n1c;
         s1 += w[D_BNW] * n1l;
         s1 += w[D_TC] * q3Bc;
         s1 += w[D_TW] * q3Bl;
         s1 += w[D_TS] * q3Ac;
         s1 += w[D_TSE] * q3Ar;

         const bool ok0 = ( yo >= yMin ) && ( yo <= yMax0 );
         const bool ok1 = ( yo >= yMin ) && ( yo <= yMax1 );
         double     e0 = 0.0, e1 = 0.0;
         if ( ModeTraits< MODE >::nExtra )
         {
            const double* er = sdata + ( j * NR + NSRC + 2 * p ) * ROWLEN + xl;
            e0               = er[( ob0 + h.x0 + cpar ) & 1];
            e1               = er[ROWLEN + ( ( ob1 + h.x0 + cpar ) & 1 )];
         }
         if ( ok0 )
            hb_epilogue_stream< MODE >( a, (size_t) ( cellBase + ob0 + x ), s0, w[D_C], invd, q1Bc, e0 );
         if ( ok1 )
            hb_epilogue_stream< MODE >( a, (size_t) ( cellBase + ob1 + x ), s1, w[D_C], invd, q2Bc, e1 );

         // slide the register window
         q0Bc = n0c;
         q0Br = n0r;
         q1Ac = q1Bc;
         q1Ar = q1Br;
         q1Bl = n1l;
         q1Bc = n1c;
         q1Br = n1r;
         q2Ac = q2Bc;
         q2Ar = q2Br;
         q2Bl = n2l;
         q2Bc = n2c;
         q2Br = n2r;
         q3Ac = q3Bc;
         q3Ar = q3Br;
         q3Bl = n3l;
         q3Bc = n3c;
         q3Br = n3r;
#pragma unroll
         for ( int q = 0; q < 4; q++ )
            gi[q] += ( N + 2 - zA - q ) - yr; // next row of slice zA-1+q
         ob0 += ( N + 1 - zA ) - yo;
         ob1 += ( N - zA ) - yo;
      }
      __syncwarp();
      if ( lane == 0 )
         mbar_arrive( &empty[stage] );
      if ( ++stage == STAGES )
      {
         stage = 0;
         phase ^= 1u;
      }
   }
}

} // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
#include <stdexcept>
#include <string>

#define HB_STREAM_XW 2 // warps along x per item: 1 -> items of 16 slices x 32 columns, 2 -> 8 slices x 64 columns

namespace {

int smCount()
{
   static int n = 0;
   if ( n == 0 )
   {
      int dev = 0;
      cudaGetDevice( &dev );
      cudaDeviceGetAttribute( &n, cudaDevAttrMultiProcessorCount, dev );
      if ( n <= 0 )
         n = 148;
   }
   return n;
}

template < int MODE, int XW, int R, int STAGES >
void launchMode( const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st )
{
   using SM           = Smem< MODE, XW, R, STAGES >;
   static bool once   = false;
   auto        kernel = hb_stream_kernel< MODE, XW, R, STAGES >;
   if ( !once )
   {
      cudaError_t e = cudaFuncSetAttribute( kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::total );
      if ( e == cudaSuccess )
         e = cudaFuncSetAttribute( kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared );
      if ( e != cudaSuccess )
         throw std::runtime_error( std::string( "hyteg_b200: cudaFuncSetAttribute: " ) + cudaGetErrorString( e ) );
      once = true;
   }
   const long long total = (long long) lv.nCells * sa.nItems;
   long long       grid  = 2LL * smCount();
   if ( grid > total )
      grid = total;
   if ( grid < 1 )
      return;
   kernel<<< (unsigned) grid, THREADS, SM::total, st >>>( lv, a, sa );
}

} // namespace

bool hb_stream_supported( int mode, const HbLevelView& lv )
{
   if ( lv.N < 32 || lv.nCells == 0 )
      return false;
   return mode == HB_APPLY_REPLACE || mode == HB_APPLY_ADD || mode == HB_RESIDUAL || mode == HB_JACOBI || mode == HB_CHEB_FIRST ||
          mode == HB_CHEB_STEP;
}

void hb_launch_stream_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st )
{
   constexpr int XW = HB_STREAM_XW;
   switch ( mode )
   {
   case HB_APPLY_REPLACE:
      launchMode< HB_APPLY_REPLACE, XW, 2, 6 >( lv, a, sa, st );
      break;
   case HB_APPLY_ADD:
      launchMode< HB_APPLY_ADD, XW, 2, 4 >( lv, a, sa, st );
      break;
   case HB_RESIDUAL:
      launchMode< HB_RESIDUAL, XW, 2, 4 >( lv, a, sa, st );
      break;
   case HB_JACOBI:
      launchMode< HB_JACOBI, XW, 2, 4 >( lv, a, sa, st );
      break;
   case HB_CHEB_FIRST:
      launchMode< HB_CHEB_FIRST, XW, 2, 4 >( lv, a, sa, st );
      break;
   case HB_CHEB_STEP:
      launchMode< HB_CHEB_STEP, XW, 2, 4 >( lv, a, sa, st );
      break;
   default:
      throw std::runtime_error( "hyteg_b200: mode not supported by the streamed kernel" );
   }
}

void hb_stream_build_items( int N, std::vector< HbItem >& items )
{
   const int ZBh = 2 * ( 8 / HB_STREAM_XW ), XTh = 32 * HB_STREAM_XW, YS = 30;
   items.clear();
   for ( int z0 = 1; z0 <= N - 3; z0 += ZBh )
      for ( int x0 = 1; x0 <= N - 2 - z0; x0 += XTh )
      {
         const int yMax = N - 1 - z0 - x0; // last row with a valid point at column x0 in slice z0
         for ( int ya = 1; ya <= yMax; ya += YS )
         {
            HbItem it;
            it.z0   = (int16_t) z0;
            it.x0   = (int16_t) x0;
            it.ya   = (int16_t) ya;
            it.nOut = (int16_t) ( ( yMax - ya + 1 < YS ) ? ( yMax - ya + 1 ) : YS );
            items.push_back( it );
         }
      }
}

#else // HB_EMUL: the streamed kernel needs TMA + mbarriers; the emulation build always uses the generic kernel

bool hb_stream_supported( int, const HbLevelView& )
{
   return false;
}
void hb_launch_stream_stencil( int, const HbLevelView&, const HbStencilArgs&, const HbStreamArgs&, cudaStream_t ) {}
void hb_stream_build_items( int, std::vector< HbItem >& items )
{
   items.clear();
}

#endif
