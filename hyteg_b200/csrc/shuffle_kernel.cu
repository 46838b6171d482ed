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
#include <type_traits>

namespace {

// CTA shape: ZP z-slice pairs x XCH column chunks = PAIRS warps.  Measured (level 8, us): 8x1 64.3, 4x1 62.0, 2x1 61.9, 4x2 68.3,
// 2x4 73.9 -- small CTAs shorten the tail of the grid, column chunks in one CTA do not pay.  More column chunks per CTA make the CTA's DRAM accesses
// longer contiguous runs per row (XCH * 240 B), fewer z pairs cost more halo slices (2 ZP + 2 loaded per 2 ZP outputs).
#ifndef HB_ZP
#define HB_ZP 4
#endif
#ifndef HB_XCH
#define HB_XCH 1
#endif
constexpr int ZP    = HB_ZP;
constexpr int XCH   = HB_XCH;
constexpr int PAIRS = ZP * XCH; // warps per CTA
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

// value stored at the output point; Ax = full stencil sum, e = fused operand (dst / rhs / x), centre = src at the point
template < int MODE >
__device__ __forceinline__ double epilogue( const HbStencilArgs& a, double Ax, double invd, double centre, double e )
{
   if ( MODE == HB_APPLY_ADD )
      return e + Ax;
   if ( MODE == HB_RESIDUAL )
      return e - Ax;
   if ( MODE == HB_JACOBI )
      return centre + a.omega * ( invd * ( e - Ax ) );
   if ( MODE == HB_CHEB_FIRST )
      return invd * ( e - Ax );
   if ( MODE == HB_SOR ) // e = rhs; Ax includes the centre term: invd * ( rhs - sum over neighbours ) = invd * ( e - Ax ) + centre
      return hb_sor_value( a.omega, Ax, invd, centre, e );
   if ( MODE == HB_CHEB_STEP ) // second store ( aux = e + omega * centre ) is issued by the caller
      return invd * Ax;
   return Ax; // HB_APPLY_REPLACE
}

} // namespace

namespace {

template < int MODE, int DEPTH >
__global__ void __maxnreg__( 128 ) hb_shuffle_kernel( const HbLevelView lv, const HbStencilArgs a, const HbStreamArgs sa )
{
   constexpr bool EX = ModeTraits< MODE >::hasExtra;
   const int      N  = lv.N;
   // item-major order: CTAs of consecutive block ids work on the same item of different macro-cells
   int       itemIdx, cl; // cl: macro-cell relative to the first one of this launch = entry of its weights behind constBase
   const int nc = sa.cellCount;
   if ( sa.cellMajor )
   {
      cl      = blockIdx.x / sa.nItems;
      itemIdx = blockIdx.x - cl * sa.nItems;
   }
   else
   {
      itemIdx = blockIdx.x / nc;
      cl      = blockIdx.x - itemIdx * nc;
   }
   const int cell = sa.cellFirst + cl;
   if ( !( ( sa.selMask >> c_bflag[sa.constBase + cl] ) & 1 ) )
      return;
   const HbItem it   = sa.items[itemIdx];
   const int    warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   const int    zA   = it.z0 + 2 * ( warp % ZP ); // this warp's output slices zA, zA+1
   if ( zA > N - 3 )
      return;
   const int x0w = it.x0 + XOUT * ( warp / ZP ); // first output column of this warp
   if ( N - 1 - zA - x0w < it.ya )
      return; // the warp's columns have no interior point in the item's rows
   const int    xc       = x0w - 1 + lane; // this lane's column
   const size_t cellBase = (size_t) cell * lv.cellSize;

   const double* w    = c_W[sa.constBase + cl];
   const double  invd = w[15]; // 1 / w[D_C], stored behind the 15 weights

   // ---- addressing ------------------------------------------------------------------------------------------------
   // Row pitch of slice z at row y is (N + 1 - z) - y.  With P pointing at (column xc, row y) of a slice whose pitch there is
   // 'pitch', row y + j of the slice 'q' slices above starts 8 * ( j * pitch - j (j - 1) / 2 - q j ) bytes further: one
   // IMAD.WIDE with an immediate multiplier per access plus an immediate offset, all accesses of an unrolled block
   // independent of each other.  Addresses are opaque 64-bit integers and the accesses are PTX, so the compiler neither
   // re-derives them from indices nor serialises them.
   using u64       = unsigned long long;
   const u64 srcB  = (u64) ( sa.srcArena + cellBase );
   const u64 dstB  = (u64) ( a.dst + cellBase );
   const u64 exB   = EX ? (u64) ( sa.extraArena + cellBase ) : 0ull;
   const int ya    = it.ya, nOut = it.nOut; // CTA-uniform trip count: keeps the shuffles in provably convergent code
   const int rowIn = row_base( N, ya - 1, zA - 1 ) + xc; // slice zA-1, row ya-1
   u64       P0    = srcB + 8ull * (u64) rowIn;
   u64       P1    = srcB + 8ull * (u64) ( row_base( N, ya - 1, zA ) + xc );
   u64       P2    = srcB + 8ull * (u64) ( row_base( N, ya - 1, zA + 1 ) + xc );
   u64       P3    = srcB + 8ull * (u64) ( row_base( N, ya - 1, zA + 2 ) + xc );
   int       ld    = ( N + 2 - zA ) - ( ya - 1 ); // pitch of slice zA-1 at the row the P's point to
   // CHAIN2 (plain apply): the partial sums travel lane x-1 -> x -> x+1 as accumulator seeds (no DADD at all), so the output of
   // column x ends up in lane x+1; the fused modes need the centre value / operand of the output column and keep it in lane x.
   constexpr bool CHAIN2 = ( MODE == HB_APPLY_REPLACE );
   const int      xo     = CHAIN2 ? xc - 1 : xc; // output column of this lane
   const int oi0 = row_base( N, ya, zA ) + xo, oi1 = row_base( N, ya, zA + 1 ) + xo;
   u64       O0 = dstB + 8ull * (u64) oi0, O1 = dstB + 8ull * (u64) oi1; // output row of the block's first iteration
   u64       E0 = exB + 8ull * (u64) oi0, E1 = exB + 8ull * (u64) oi1;   // next row of the fused operand to load
   int       so = ( N + 1 - zA ) - ya;                                   // pitch of slice zA at the row O0 / E0 point to
   int       se = so;
   const long long auxOff = ( MODE == HB_CHEB_STEP ) ? (long long) ( (u64) a.aux - (u64) a.dst ) : 0ll;

   const bool laneOut = CHAIN2 ? ( lane >= 2 ) : ( ( lane >= 1 ) && ( lane <= XOUT ) );
   const int  yMax0   = laneOut ? min( ya + nOut - 1, min( N - 2 - zA, N - 1 - zA - xo ) ) : -1;
   const int  yMax1   = ( laneOut && zA + 1 <= N - 3 ) ? min( ya + nOut - 1, min( N - 3 - zA, N - 2 - zA - xo ) ) : -1;
   int        r0 = max( 0, yMax0 - ya + 1 ), r1 = max( 0, yMax1 - ya + 1 ); // rows this lane still has to store in slice zA / zA+1
   // SOR: colour (x + 2y + 3z) mod 4 of the output point (same colouring as the generic kernel); k0 == 0 mod 4 <=> slice zA matches
   int k0 = ( MODE == HB_SOR ) ? ( xc + 2 * ya + 3 * zA - a.colour ) : 0;

   auto rowAddr = [&]( u64 P, int pitch, int j, int q ) -> u64 { return P + (u64) ( (long long) pitch * ( 8 * j ) + (long long) ( -8 * ( j * ( j - 1 ) / 2 + q * j ) ) ); };
   auto load    = [&]( double& v, u64 addr, bool pred ) {
      if ( MODE == HB_SOR ) // SOR updates the array it reads (only points of other colours matter): coherent loads
         asm volatile( "{\n .reg .pred p;\n setp.ne.s32 p, %2, 0;\n @p ld.global.f64 %0, [%1];\n}" : "+d"( v ) : "l"( addr ), "r"( (int) pred ) : "memory" );
      else
         asm( "{\n .reg .pred p;\n setp.ne.s32 p, %2, 0;\n @p ld.global.nc.f64 %0, [%1];\n}" : "+d"( v ) : "l"( addr ), "r"( (int) pred ) );
   };
   auto loadNow = [&]( double& v, u64 addr ) {
      if ( MODE == HB_SOR )
         asm volatile( "ld.global.f64 %0, [%1];" : "=d"( v ) : "l"( addr ) : "memory" );
      else
         asm( "ld.global.nc.f64 %0, [%1];" : "=d"( v ) : "l"( addr ) );
   };
   auto store = [&]( u64 addr, double v, bool pred ) {
      asm volatile( "{\n .reg .pred p;\n setp.ne.s32 p, %2, 0;\n @p st.global.f64 [%0], %1;\n}" ::"l"( addr ), "d"( v ), "r"( (int) pred ) : "memory" );
   };

   // ---- prologue: window rows ya-1 (A) and ya (B), then the ring: rows ya+1 .. ya+DEPTH (and the fused operand of rows ya ..) ----
   double B0, A1, B1, A2, B2, A3, B3;
   loadNow( A1, rowAddr( P1, ld, 0, 1 ) );
   loadNow( A2, rowAddr( P2, ld, 0, 2 ) );
   loadNow( A3, rowAddr( P3, ld, 0, 3 ) );
   loadNow( B0, rowAddr( P0, ld, 1, 0 ) );
   loadNow( B1, rowAddr( P1, ld, 1, 1 ) );
   loadNow( B2, rowAddr( P2, ld, 1, 2 ) );
   loadNow( B3, rowAddr( P3, ld, 1, 3 ) );
   double pf[DEPTH][4];
   double pe[DEPTH][2];
   int    nLoad = nOut; // ring rows still to load: ya+1 .. ya+nOut
#pragma unroll
   for ( int d = 0; d < DEPTH; d++ )
   {
      pf[d][0] = pf[d][1] = pf[d][2] = pf[d][3] = 0.0;
      pe[d][0] = pe[d][1] = 0.0;
      const bool ok = d < nLoad;
      load( pf[d][0], rowAddr( P0, ld, 2 + d, 0 ), ok );
      load( pf[d][1], rowAddr( P1, ld, 2 + d, 1 ), ok );
      load( pf[d][2], rowAddr( P2, ld, 2 + d, 2 ), ok );
      load( pf[d][3], rowAddr( P3, ld, 2 + d, 3 ), ok );
      if ( EX )
      {
         load( pe[d][0], rowAddr( E0, se, d, 0 ), ok );
         load( pe[d][1], rowAddr( E1, se, d, 1 ), ok );
      }
   }
   P0 = rowAddr( P0, ld, 2 + DEPTH, 0 );
   P1 = rowAddr( P1, ld, 2 + DEPTH, 1 );
   P2 = rowAddr( P2, ld, 2 + DEPTH, 2 );
   P3 = rowAddr( P3, ld, 2 + DEPTH, 3 );
   ld -= 2 + DEPTH;
   if ( EX )
   {
      E0 = rowAddr( E0, se, DEPTH, 0 );
      E1 = rowAddr( E1, se, DEPTH, 1 );
      se -= DEPTH;
   }
   nLoad -= DEPTH;

   // one unrolled block of DEPTH output rows; guarded = the refill loads test nLoad (tail of the item), otherwise unconditional
   auto block = [&]( auto guarded ) {
      constexpr bool G  = decltype( guarded )::value;
      const int      km = k0 & 3;
#pragma unroll
      for ( int u = 0; u < DEPTH; u++ )
      {
         const double N0 = pf[u][0], N1 = pf[u][1], N2 = pf[u][2], N3 = pf[u][3];
         const double x0 = pe[u][0], x1 = pe[u][1];
         // refill the slot: row DEPTH ahead
         if ( G )
         {
            const bool ok = u < nLoad;
            load( pf[u][0], rowAddr( P0, ld, u, 0 ), ok );
            load( pf[u][1], rowAddr( P1, ld, u, 1 ), ok );
            load( pf[u][2], rowAddr( P2, ld, u, 2 ), ok );
            load( pf[u][3], rowAddr( P3, ld, u, 3 ), ok );
            if ( EX )
            {
               load( pe[u][0], rowAddr( E0, se, u, 0 ), ok );
               load( pe[u][1], rowAddr( E1, se, u, 1 ), ok );
            }
         }
         else
         {
            loadNow( pf[u][0], rowAddr( P0, ld, u, 0 ) );
            loadNow( pf[u][1], rowAddr( P1, ld, u, 1 ) );
            loadNow( pf[u][2], rowAddr( P2, ld, u, 2 ) );
            loadNow( pf[u][3], rowAddr( P3, ld, u, 3 ) );
            if ( EX )
            {
               loadNow( pe[u][0], rowAddr( E0, se, u, 0 ) );
               loadNow( pe[u][1], rowAddr( E1, se, u, 1 ) );
            }
         }

         // left 4-tap partial sums of both outputs: evaluated for the lane to the right
         double l0 = w[D_W] * B1;
         l0        = fma( w[D_NW], N1, l0 );
         l0        = fma( w[D_BNW], N0, l0 );
         l0        = fma( w[D_TW], B2, l0 );
         double l1 = w[D_W] * B2;
         l1        = fma( w[D_NW], N2, l1 );
         l1        = fma( w[D_BNW], N1, l1 );
         l1        = fma( w[D_TW], B3, l1 );
         // centre-column taps, seeded with the left neighbour's partial sum.  output (zA, y): below = slice 0, own = 1, above = 2
         double c0 = __shfl_up_sync( 0xffffffffu, l0, 1 );
         c0        = fma( w[D_C], B1, c0 );
         c0        = fma( w[D_N], N1, c0 );
         c0        = fma( w[D_S], A1, c0 );
         c0        = fma( w[D_BC], B0, c0 );
         c0        = fma( w[D_BN], N0, c0 );
         c0        = fma( w[D_TC], B2, c0 );
         c0        = fma( w[D_TS], A2, c0 );
         // output (zA+1, y): below = slice 1, own = 2, above = 3
         double c1 = __shfl_up_sync( 0xffffffffu, l1, 1 );
         c1        = fma( w[D_C], B2, c1 );
         c1        = fma( w[D_N], N2, c1 );
         c1        = fma( w[D_S], A2, c1 );
         c1        = fma( w[D_BC], B1, c1 );
         c1        = fma( w[D_BN], N1, c1 );
         c1        = fma( w[D_TC], B3, c1 );
         c1        = fma( w[D_TS], A3, c1 );
         double Ax0, Ax1;
         if ( CHAIN2 )
         {
            // right taps: this lane's values are the right neighbours of column x-1, whose sum so far arrives from lane x-1
            Ax0 = __shfl_up_sync( 0xffffffffu, c0, 1 );
            Ax0 = fma( w[D_E], B1, Ax0 );
            Ax0 = fma( w[D_SE], A1, Ax0 );
            Ax0 = fma( w[D_BE], B0, Ax0 );
            Ax0 = fma( w[D_TSE], A2, Ax0 );
            Ax1 = __shfl_up_sync( 0xffffffffu, c1, 1 );
            Ax1 = fma( w[D_E], B2, Ax1 );
            Ax1 = fma( w[D_SE], A2, Ax1 );
            Ax1 = fma( w[D_BE], B1, Ax1 );
            Ax1 = fma( w[D_TSE], A3, Ax1 );
         }
         else
         {
            double r0v = w[D_E] * B1;
            r0v        = fma( w[D_SE], A1, r0v );
            r0v        = fma( w[D_BE], B0, r0v );
            r0v        = fma( w[D_TSE], A2, r0v );
            double r1v = w[D_E] * B2;
            r1v        = fma( w[D_SE], A2, r1v );
            r1v        = fma( w[D_BE], B1, r1v );
            r1v        = fma( w[D_TSE], A3, r1v );
            Ax0        = c0 + __shfl_down_sync( 0xffffffffu, r0v, 1 );
            Ax1        = c1 + __shfl_down_sync( 0xffffffffu, r1v, 1 );
         }

         bool s0 = u < r0, s1 = u < r1;
         if ( MODE == HB_SOR )
         {
            s0 = s0 && ( km == ( ( 4 - 2 * u ) & 3 ) );
            s1 = s1 && ( km == ( ( 1 - 2 * u ) & 3 ) ); // k0 + 3 + 2u == 0 mod 4
         }
         const u64 q0 = rowAddr( O0, so, u, 0 ), q1 = rowAddr( O1, so, u, 1 );
         store( q0, epilogue< MODE >( a, Ax0, invd, B1, x0 ), s0 );
         store( q1, epilogue< MODE >( a, Ax1, invd, B2, x1 ), s1 );
         if ( MODE == HB_CHEB_STEP )
         {
            store( q0 + (u64) auxOff, x0 + a.omega * B1, s0 );
            store( q1 + (u64) auxOff, x1 + a.omega * B2, s1 );
         }

         // slide the window
         B0 = N0;
         A1 = B1;
         B1 = N1;
         A2 = B2;
         B2 = N2;
         A3 = B3;
         B3 = N3;
      }
      P0 = rowAddr( P0, ld, DEPTH, 0 );
      P1 = rowAddr( P1, ld, DEPTH, 1 );
      P2 = rowAddr( P2, ld, DEPTH, 2 );
      P3 = rowAddr( P3, ld, DEPTH, 3 );
      ld -= DEPTH;
      O0 = rowAddr( O0, so, DEPTH, 0 );
      O1 = rowAddr( O1, so, DEPTH, 1 );
      so -= DEPTH;
      if ( EX )
      {
         E0 = rowAddr( E0, se, DEPTH, 0 );
         E1 = rowAddr( E1, se, DEPTH, 1 );
         se -= DEPTH;
      }
      r0 -= DEPTH;
      r1 -= DEPTH;
      nLoad -= DEPTH;
      if ( MODE == HB_SOR )
         k0 += 2 * DEPTH;
   };
   int left = nOut; // output rows still to produce
   while ( nLoad >= DEPTH )
   {
      asm volatile( "" ::: "memory" ); // keeps the unrolled blocks apart: without it ptxas schedules the fused modes 7 % longer
      block( std::false_type{} );
      left -= DEPTH;
   }
   for ( ; left > 0; left -= DEPTH )
      block( std::true_type{} );
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
   if ( lv.N < 32 || lv.nCells == 0 )
      return false;
   return mode == HB_APPLY_REPLACE || mode == HB_APPLY_ADD || mode == HB_RESIDUAL || mode == HB_JACOBI || mode == HB_CHEB_FIRST ||
          mode == HB_CHEB_STEP || mode == HB_SOR;
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
   case HB_SOR:
      hb_shuffle_kernel< HB_SOR, DEPTH ><<< g, b, 0, st >>>( lv, a, sa );
      break;
   default:
      throw std::runtime_error( "hyteg_b200: mode not supported by the shuffle kernel" );
   }
}

void hb_launch_shuffle_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st, int depth )
{
   const long long grid = (long long) sa.cellCount * sa.nItems;
   if ( grid < 1 )
      return;
   const unsigned g = (unsigned) grid, b = PAIRS * 32;
   if ( depth <= 0 ) // measured best on B200 with 4-warp CTAs (profiles/r1_b_interior_sweeps.txt)
      depth = 4;
   switch ( depth )
   {
   case 3:
      launchDepth< 3 >( mode, g, b, lv, a, sa, st );
      break;
   case 4:
      launchDepth< 4 >( mode, g, b, lv, a, sa, st );
      break;
   case 7:
      launchDepth< 7 >( mode, g, b, lv, a, sa, st );
      break;
   default:
      launchDepth< DEPTH_DEFAULT >( mode, g, b, lv, a, sa, st );
   }
}

void hb_shuffle_build_items( int N, std::vector< HbItem >& items, int ys, int order )
{
   // ys + 2 input rows per item: a multiple of DEPTH_DEFAULT keeps the last unrolled block full
   const int ZBh = 2 * ZP, YS = ys > 0 ? ys : 32;
   items.clear();
   std::vector< long long > work;
   for ( int z0 = 1; z0 <= N - 3; z0 += ZBh )
      for ( int x0 = 1; x0 <= N - 2 - z0; x0 += XOUT * XCH )
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
      const int nz = std::min( 2 * ZP, N - 2 - it.z0 );
      const int nx = std::min( XCH, ( N - 2 - it.z0 - it.x0 - it.ya ) / XOUT + 1 );
      return (long long) ( it.nOut + 2 + DEPTH_DEFAULT ) * ( ( nz + 1 ) / 2 ) * nx;
   };
   if ( order == 2 ) // x fastest: consecutive CTAs read neighbouring column chunks of the same rows
      std::stable_sort( items.begin(), items.end(), []( const HbItem& a, const HbItem& b ) {
         return a.z0 != b.z0 ? a.z0 < b.z0 : ( a.ya != b.ya ? a.ya < b.ya : a.x0 < b.x0 );
      } );
   if ( order == 0 )
      std::stable_sort( items.begin(), items.end(), [&]( const HbItem& a, const HbItem& b ) { return cost( a ) > cost( b ); } );
}

#else // HB_EMUL

bool hb_shuffle_supported( int, const HbLevelView& )
{
   return false;
}
void hb_launch_shuffle_stencil( int, const HbLevelView&, const HbStencilArgs&, const HbStreamArgs&, cudaStream_t, int ) {}
void hb_shuffle_build_items( int, std::vector< HbItem >& items, int, int )
{
   items.clear();
}
void hb_shuffle_upload_weights( int, int, const double*, const uint8_t*, cudaStream_t ) {}

#endif
