// Interior stencil kernel, "centre-column + shuffle" form (sm_100a): the hot loop of apply / residual / Jacobi / Chebyshev.
//
// Replaces apply_3D_macrocell_vertexdof_to_vertexdof_{replace,add} (reference
// src/constant_stencil_operator/P1generatedKernels/apply_3D_macrocell_vertexdof_to_vertexdof_replace.cpp:52-77)
// and the BLAS-1 passes the reference chains behind it (P1Operator.hpp:429-447, ChebyshevSmoother.hpp:179-222).
//
// Why this form (measurements in profiles/, history in DESIGN.md section 2):
//   * r1_a, per-point l/c/r loads through L1: 9 loads + ~200 thread instructions per point, latency/issue bound;
//   * r1_b..e, cp.async.bulk (1D TMA) row segments into a shared-memory ring: correct and bit-identical, but a row segment
//     of the tetrahedral layout is <= 544 B and the TMA unit serves about one bulk request per ~46 cycles per SM
//     -> 6-8 B/clk/SM, consumers starve.  Rows have a different pitch in every row, so no tensor map can describe a slab.
// Here every thread owns ONE column x and TWO z slices and marches along y.  It loads only the centre-column value of each
// new row of its four input slices (4 coalesced LDG.64 per 2 outputs) and splits the 15-point stencil by column:
//     out(x) = C(x) + L(x-1) + R(x+1),   C = 7 centre-column taps, L = the 4 taps at x-1, R = the 4 taps at x+1,
// where L and R are evaluated by the neighbouring lanes from THEIR centre values and handed over with one shuffle each.
// The register window is 11 doubles instead of 28, which leaves room for a 3-row-deep register prefetch ring: the loads
// of row y+3 are in flight while row y is computed, no shared memory, no barriers, no producer warps.
// A warp covers 32 columns and produces 30 outputs per row (lanes 0 and 31 are halo lanes).
#include "shuffle_kernel.h"

#ifndef HB_EMUL

#include <cstdint>

namespace {

constexpr int PAIRS = 8;  // warps per CTA = z-slice pairs per work item -> 16 output slices
constexpr int XOUT  = 30; // output columns per warp
constexpr int DEPTH_DEFAULT = 5; // rows in flight ahead of the row being computed (template parameter DEPTH)

// Interior stencil weights of (level, macro-cell) pairs.  Read with a CTA-uniform index, so they live in UNIFORM registers
// (LDCU.64 UR, c[0x3][UR+imm]) and enter the DFMAs as uniform operands: no vector registers are spent on the 15 weights,
// which is what makes room for the DEPTH-row prefetch ring at 2 x 8 warps per SM.
__constant__ double  c_W[HB_CONST_W_ENTRIES][16];
__constant__ uint8_t c_bflag[HB_CONST_W_ENTRIES]; // meshBoundaryFlag of the macro-cell of each entry

__device__ __forceinline__ int tet( int k )
{
   return ( k * ( k + 1 ) * ( k + 2 ) ) / 6;
}
// index of (x=0, y, z) inside a macro-cell array (reference indexing/MacroCellIndexing.hpp:44-52)
__device__ __forceinline__ int row_base( int N, int y, int z )
{
   const int wz = N + 1 - z;
   return tet( N + 1 ) - tet( wz ) + y * ( wz + 1 ) - ( y * ( y + 1 ) ) / 2;
}

template < int MODE >
struct ModeTraits
{
   static constexpr bool hasExtra = ( MODE != HB_APPLY_REPLACE );
};

template < int MODE >
__device__ __forceinline__ void epilogue( const HbStencilArgs& a, size_t gi, double Ax, double invd, double centre, double e )
{
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

} // namespace

namespace {

template < int MODE, int DEPTH >
__global__ void __launch_bounds__( PAIRS * 32, 2 ) hb_shuffle_kernel( const HbLevelView lv, const HbStencilArgs a, const HbStreamArgs sa )
{
   constexpr bool EX = ModeTraits< MODE >::hasExtra;
   const int      N  = lv.N;
   // item-major order, items sorted by decreasing work: the last wave of CTAs consists of the smallest items
   const int itemIdx = blockIdx.x / lv.nCells;
   const int cell    = blockIdx.x - itemIdx * lv.nCells;
   if ( !( ( sa.selMask >> c_bflag[sa.constBase + cell] ) & 1 ) )
      return;
   const HbItem it   = sa.items[itemIdx];
   const int    warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   const int    zA   = it.z0 + 2 * warp; // this warp's output slices zA, zA+1
   if ( zA > N - 3 )
      return;
   const int    xc       = it.x0 - 1 + lane; // this lane's column
   const size_t cellBase = (size_t) cell * lv.cellSize;
   const double* __restrict__ S = sa.srcArena + cellBase;
   const double* __restrict__ E = EX ? sa.extraArena + cellBase : nullptr;

   const double* w    = c_W[sa.constBase + cell];
   const double  invd = w[15]; // 1 / w[D_C], stored behind the 15 weights

   const int yFirstIn = it.ya - 1;       // first input row
   const int yLastIn  = it.ya + it.nOut; // last input row
   // running indices: li[q] = (column xc, next row to LOAD) of slice zA-1+q; ei/oi = extra operand / output position
   int li[4];
#pragma unroll
   for ( int q = 0; q < 4; q++ )
      li[q] = row_base( N, yFirstIn, zA - 1 + q ) + xc;
   int       yLoad = yFirstIn;
   int       oi0 = row_base( N, yFirstIn - 1, zA ) + xc, oi1 = row_base( N, yFirstIn - 1, zA + 1 ) + xc; // output row of iteration 0
   int       ei0 = oi0, ei1 = oi1;                                                                       // extra row to LOAD
   int       yExtra = yFirstIn - 1;
   const bool laneOut = ( lane >= 1 ) && ( lane <= XOUT );
   const int  yMax0   = laneOut ? min( it.ya + it.nOut - 1, min( N - 2 - zA, N - 1 - zA - xc ) ) : -1;
   const int  yMax1   = ( laneOut && zA + 1 <= N - 3 ) ? min( it.ya + it.nOut - 1, min( N - 3 - zA, N - 2 - zA - xc ) ) : -1;

   double pf[DEPTH][4];
   double pe[DEPTH][2];
   auto   loadRow = [&]( double ( &dst )[4], double ( &de )[2] ) {
      const bool ok = yLoad <= yLastIn;
#pragma unroll
      for ( int q = 0; q < 4; q++ )
      {
         dst[q] = ok ? __ldg( S + li[q] ) : 0.0;
         li[q] += ( N + 2 - zA - q ) - yLoad; // next row of slice zA-1+q
      }
      if ( EX )
      {
         const bool oke = ( yExtra >= it.ya ) && ( yExtra < yLastIn );
         de[0]          = ( oke && yExtra <= yMax0 ) ? E[ei0] : 0.0;
         de[1]          = ( oke && yExtra <= yMax1 ) ? E[ei1] : 0.0;
         ei0 += ( N + 1 - zA ) - yExtra;
         ei1 += ( N - zA ) - yExtra;
      }
      yLoad++;
      yExtra++;
   };
#pragma unroll
   for ( int d = 0; d < DEPTH; d++ )
      loadRow( pf[d], pe[d] );

   // register window (centre column only): slice q rows A (y-1), B (y)
   double B0 = 0, A1 = 0, B1 = 0, A2 = 0, B2 = 0, A3 = 0, B3 = 0;
   int    yo = yFirstIn - 1; // output row of the current iteration
   const int nIn = it.nOut + 2;
   for ( int base = 0; base < nIn; base += DEPTH )
   {
#pragma unroll
      for ( int u = 0; u < DEPTH; u++ )
      {
         const double N0 = pf[u][0], N1 = pf[u][1], N2 = pf[u][2], N3 = pf[u][3];
         const double e0 = pe[u][0], e1 = pe[u][1];
         loadRow( pf[u], pe[u] ); // refill the slot: row DEPTH ahead

         // output (zA, yo): below = slice 0, own = 1, above = 2
         double c0 = w[D_C] * B1;
         c0        = fma( w[D_N], N1, c0 );
         c0        = fma( w[D_S], A1, c0 );
         c0        = fma( w[D_BC], B0, c0 );
         c0        = fma( w[D_BN], N0, c0 );
         c0        = fma( w[D_TC], B2, c0 );
         c0        = fma( w[D_TS], A2, c0 );
         double l0 = w[D_W] * B1;
         l0        = fma( w[D_NW], N1, l0 );
         l0        = fma( w[D_BNW], N0, l0 );
         l0        = fma( w[D_TW], B2, l0 );
         double r0 = w[D_E] * B1;
         r0        = fma( w[D_SE], A1, r0 );
         r0        = fma( w[D_BE], B0, r0 );
         r0        = fma( w[D_TSE], A2, r0 );
         // output (zA+1, yo): below = slice 1, own = 2, above = 3
         double c1 = w[D_C] * B2;
         c1        = fma( w[D_N], N2, c1 );
         c1        = fma( w[D_S], A2, c1 );
         c1        = fma( w[D_BC], B1, c1 );
         c1        = fma( w[D_BN], N1, c1 );
         c1        = fma( w[D_TC], B3, c1 );
         c1        = fma( w[D_TS], A3, c1 );
         double l1 = w[D_W] * B2;
         l1        = fma( w[D_NW], N2, l1 );
         l1        = fma( w[D_BNW], N1, l1 );
         l1        = fma( w[D_TW], B3, l1 );
         double r1 = w[D_E] * B2;
         r1        = fma( w[D_SE], A2, r1 );
         r1        = fma( w[D_BE], B1, r1 );
         r1        = fma( w[D_TSE], A3, r1 );

         const double Ax0 = c0 + __shfl_up_sync( 0xffffffffu, l0, 1 ) + __shfl_down_sync( 0xffffffffu, r0, 1 );
         const double Ax1 = c1 + __shfl_up_sync( 0xffffffffu, l1, 1 ) + __shfl_down_sync( 0xffffffffu, r1, 1 );

         if ( yo >= it.ya && yo <= yMax0 )
            epilogue< MODE >( a, cellBase + oi0, Ax0, invd, B1, e0 );
         if ( yo >= it.ya && yo <= yMax1 )
            epilogue< MODE >( a, cellBase + oi1, Ax1, invd, B2, e1 );

         // slide the window
         B0 = N0;
         A1 = B1;
         B1 = N1;
         A2 = B2;
         B2 = N2;
         A3 = B3;
         B3 = N3;
         oi0 += ( N + 1 - zA ) - yo;
         oi1 += ( N - zA ) - yo;
         yo++;
      }
   }
}

} // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
#include <algorithm>
#include <stdexcept>
#include <string>

void hb_shuffle_upload_weights( int first, int n, const double* w16, const uint8_t* bflag, cudaStream_t st )
{
   if ( first < 0 || n < 0 || first + n > HB_CONST_W_ENTRIES )
      throw std::runtime_error( "hyteg_b200: constant-memory weight table overflow" );
   if ( n == 0 )
      return;
   cudaError_t e = cudaMemcpyToSymbolAsync( c_W, w16, (size_t) n * 16 * sizeof( double ), (size_t) first * 16 * sizeof( double ), cudaMemcpyHostToDevice, st );
   if ( e == cudaSuccess )
      e = cudaMemcpyToSymbolAsync( c_bflag, bflag, (size_t) n, (size_t) first, cudaMemcpyHostToDevice, st );
   if ( e != cudaSuccess )
      throw std::runtime_error( std::string( "hyteg_b200: cudaMemcpyToSymbolAsync: " ) + cudaGetErrorString( e ) );
}

bool hb_shuffle_supported( int mode, const HbLevelView& lv )
{
   if ( lv.N < 32 || lv.nCells == 0 || lv.nCells > HB_CONST_W_ENTRIES )
      return false;
   return mode == HB_APPLY_REPLACE || mode == HB_APPLY_ADD || mode == HB_RESIDUAL || mode == HB_JACOBI || mode == HB_CHEB_FIRST ||
          mode == HB_CHEB_STEP;
}

template < int DEPTH >
static void launchDepth( int mode, unsigned g, unsigned b, const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st )
{
   switch ( mode )
   {
   case HB_APPLY_REPLACE:
      hb_shuffle_kernel< HB_APPLY_REPLACE, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   case HB_APPLY_ADD:
      hb_shuffle_kernel< HB_APPLY_ADD, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   case HB_RESIDUAL:
      hb_shuffle_kernel< HB_RESIDUAL, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   case HB_JACOBI:
      hb_shuffle_kernel< HB_JACOBI, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   case HB_CHEB_FIRST:
      hb_shuffle_kernel< HB_CHEB_FIRST, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   case HB_CHEB_STEP:
      hb_shuffle_kernel< HB_CHEB_STEP, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   default:
      throw std::runtime_error( "hyteg_b200: mode not supported by the shuffle kernel" );
   }
}

void hb_launch_shuffle_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st, int depth )
{
   const long long grid = (long long) lv.nCells * sa.nItems;
   if ( grid < 1 )
      return;
   const unsigned g = (unsigned) grid, b = PAIRS * 32;
   switch ( depth )
   {
   case 3:
      launchDepth< 3 >( mode, g, b, lv, a, sa, st );
      break;
   case 7:
      launchDepth< 7 >( mode, g, b, lv, a, sa, st );
      break;
   case 9:
      launchDepth< 9 >( mode, g, b, lv, a, sa, st );
      break;
   default:
      launchDepth< DEPTH_DEFAULT >( mode, g, b, lv, a, sa, st );
   }
}

void hb_shuffle_build_items( int N, std::vector< HbItem >& items )
{
   const int ZBh = 2 * PAIRS, YS = 45;
   items.clear();
   std::vector< long long > work;
   for ( int z0 = 1; z0 <= N - 3; z0 += ZBh )
      for ( int x0 = 1; x0 <= N - 2 - z0; x0 += XOUT )
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
   // largest items first (work ~ rows x active warps): the tail of the grid is made of the small items at the diagonal
   auto cost = [N]( const HbItem& it ) {
      const int nz = std::min( 2 * PAIRS, N - 2 - it.z0 );
      return (long long) ( it.nOut + 2 + DEPTH_DEFAULT ) * ( ( nz + 1 ) / 2 );
   };
   std::stable_sort( items.begin(), items.end(), [&]( const HbItem& a, const HbItem& b ) { return cost( a ) > cost( b ); } );
}

#else // HB_EMUL

bool hb_shuffle_supported( int, const HbLevelView& )
{
   return false;
}
void hb_launch_shuffle_stencil( int, const HbLevelView&, const HbStencilArgs&, const HbStreamArgs&, cudaStream_t, int ) {}
void hb_shuffle_build_items( int, std::vector< HbItem >& items )
{
   items.clear();
}
void hb_shuffle_upload_weights( int, int, const double*, const uint8_t*, cudaStream_t ) {}

#endif
