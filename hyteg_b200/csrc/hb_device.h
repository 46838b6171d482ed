// Device-side building blocks of the P1 constant-stencil path (sm_100a; see hb_compat.h for the
// test-only host emulation).
#pragma once
#include "hb_compat.h"
#include "hb_types.h"

// ---------------------------------------------------------------------------------------------
// 15-point stencil directions (logical micro-vertex offsets), same set as HyTeG's
// stencilDirection for vertex DoFs (reference src/hyteg/p1functionspace/VertexDoFIndexing.cpp:484-525).
// index: 0 C | 1 W 2 E 3 N 4 S 5 NW 6 SE | 7 TC 8 TW 9 TS 10 TSE | 11 BC 12 BN 13 BE 14 BNW
// ---------------------------------------------------------------------------------------------
enum
{
   D_C = 0,
   D_W,
   D_E,
   D_N,
   D_S,
   D_NW,
   D_SE,
   D_TC,
   D_TW,
   D_TS,
   D_TSE,
   D_BC,
   D_BN,
   D_BE,
   D_BNW
};

HB_CONSTANT int8_t hb_dir[HB_NDIR][3] = { { 0, 0, 0 },
                                          { -1, 0, 0 },
                                          { 1, 0, 0 },
                                          { 0, 1, 0 },
                                          { 0, -1, 0 },
                                          { -1, 1, 0 },
                                          { 1, -1, 0 },
                                          { 0, 0, 1 },
                                          { -1, 0, 1 },
                                          { 0, -1, 1 },
                                          { 1, -1, 1 },
                                          { 0, 0, -1 },
                                          { 0, 1, -1 },
                                          { 1, 0, -1 },
                                          { -1, 1, -1 } };

// hb_valid[mask] : bit d set <=> (p + dir d) stays inside the closed macro-tet for a point p with
// boundary mask 'mask'.  Generated from hb_dir (see hb_make_valid() in stencil_assembly.cpp; the
// unit tests check the table against that function).
HB_CONSTANT uint16_t hb_valid[HB_NMASK] = { 0x7fff, 0x3edd, 0x79af, 0x388d, 0x07ff, 0x06dd, 0x01af, 0x008d,
                                            0x7b73, 0x3a51, 0x7923, 0x3801, 0x0373, 0x0251, 0x0123, 0x0001 };

__host__ __device__ __forceinline__ int hb_tet( int k )
{
   return ( k * ( k + 1 ) * ( k + 2 ) ) / 6;
}

// linear index of micro-vertex (x,y,z) in a macro-cell array with N subdivisions
// (reference src/hyteg/indexing/MacroCellIndexing.hpp:44-52, width = N+1)
__host__ __device__ __forceinline__ int hb_index( int N, int x, int y, int z )
{
   const int wz = N + 1 - z;
   return hb_tet( N + 1 ) - hb_tet( wz ) + y * ( wz + 1 ) - ( y * ( y + 1 ) ) / 2 + x;
}

__host__ __device__ __forceinline__ int hb_mask( int N, int x, int y, int z )
{
   return ( x == 0 ? 1 : 0 ) | ( y == 0 ? 2 : 0 ) | ( z == 0 ? 4 : 0 ) | ( x + y + z == N ? 8 : 0 );
}

// kernel modes ---------------------------------------------------------------------------------
enum HbMode
{
   HB_APPLY_REPLACE = 0, // dst  = A src
   HB_APPLY_ADD,         // dst += A src
   HB_RESIDUAL,          // dst  = rhs - A src
   HB_JACOBI,            // dst  = src + omega D^-1 (rhs - A src)
   HB_SOR,               // dst(=src) in place, one colour: (1-w) x + w/d (rhs - sum_nb)
   HB_CHEB_FIRST,        // dst  = D^-1 (rhs - A src)
   HB_CHEB_STEP,         // dst  = D^-1 A src ; aux += omega * src
   HB_DIAG,              // dst  = diag(A)
   HB_INVDIAG            // dst  = 1/diag(A)
};

struct HbStencilArgs
{
   const double* const* src; // cell pointer table (owned + mirrored cells) used for stencil reads
   const double*        srcBase; // arena of the source function: owned cell c starts at srcBase + c * cellSize
   const double*        W;   // [nSlots][16][15]
   double*              dst; // local arena
   const double*        rhs; // local arena
   double*              aux; // local arena
   double*              scratch; // [nCells*nBnd] staging for interface phases (SOR)
   double               omega;
   int32_t              flag;    // DoFType selection
   uint32_t             selMask; // bit f set: primitives with meshBoundaryFlag f are selected ( bc.type[f] & flag )
   int32_t              colour;  // SOR colour 0..3 (interior)
   int32_t              dimSel;  // boundary phase: bitmask of primitive dims to process (1<<dim)
   int32_t              toScratch; // boundary results go to scratch instead of dst
   int32_t              bndColour; // interface colour to process (-1: all), see hb_interface_colour
   HbBC                 bc;
};

template < int MODE >
__device__ __forceinline__ void
    hb_epilogue( const HbStencilArgs& a, double* out, size_t gi, double sNb, double diag, double centre )
{
   // sNb: sum over the 14 neighbours, diag: centre weight, centre: src value at the point
   if ( MODE == HB_APPLY_REPLACE )
      *out = sNb + diag * centre;
   else if ( MODE == HB_APPLY_ADD )
      *out = a.dst[gi] + ( sNb + diag * centre );
   else if ( MODE == HB_RESIDUAL )
      *out = a.rhs[gi] - ( sNb + diag * centre );
   else if ( MODE == HB_JACOBI )
      *out = centre + a.omega * ( ( 1.0 / diag ) * ( a.rhs[gi] - ( sNb + diag * centre ) ) );
   else if ( MODE == HB_SOR )
      *out = ( 1.0 - a.omega ) * centre + a.omega * ( 1.0 / diag ) * ( a.rhs[gi] - sNb );
   else if ( MODE == HB_CHEB_FIRST )
      *out = ( 1.0 / diag ) * ( a.rhs[gi] - ( sNb + diag * centre ) );
   else if ( MODE == HB_CHEB_STEP )
   {
      *out      = ( 1.0 / diag ) * ( sNb + diag * centre );
      a.aux[gi] = a.aux[gi] + a.omega * centre;
   }
   else if ( MODE == HB_DIAG )
      *out = diag;
   else if ( MODE == HB_INVDIAG )
      *out = 1.0 / diag;
}

// linear index, inside macro-cell rc.slot, of the image of the interface point (x,y,z) of the macro-cell the copy list belongs to
__host__ __device__ __forceinline__ int hb_copy_index( int N, const HbRegionCopy& rc, int x, int y, int z )
{
   const int xp = rc.c[0][0] * N + rc.c[0][1] * x + rc.c[0][2] * y + rc.c[0][3] * z;
   const int yp = rc.c[1][0] * N + rc.c[1][1] * x + rc.c[1][2] * y + rc.c[1][3] * z;
   const int zp = rc.c[2][0] * N + rc.c[2][1] * x + rc.c[2][2] * y + rc.c[2][3] * z;
   return hb_index( N, xp, yp, zp );
}

// Stencil at an interface point (x,y,z) of region 'mask' of macro-cell 'cell': sum of the partial stencils of every
// macro-cell that contains the point, visited in ascending global cell id so that all copies of the point produce
// bit-identical results.  The partial stencil of one cell is exactly HyTeG's calculateStencilInMacroCellForm_new
// (reference P1Elements.hpp:465-515); the sum over the neighbouring cells is how HyTeG builds its macro-face/edge/vertex
// stencils (reference P1Elements.hpp:626-720, 870-1017).  All threads of a block work on the same (cell, region), so the
// region-copy entries, their weights and the valid-direction masks are block-uniform.
__device__ __forceinline__ void hb_boundary_stencil_range( const HbLevelView&   lv,
                                                           const double*        W,
                                                           const double* const* src,
                                                           const double*        srcBase,
                                                           int                  e0,
                                                           int                  e1,
                                                           int                  x,
                                                           int                  y,
                                                           int                  z,
                                                           double&              sNb,
                                                           double&              diag )
{
   const int N = lv.N;
   sNb         = 0.0;
   diag        = 0.0;
   for ( int e = e0; e < e1; e++ )
   {
      const HbRegionCopy rc = lv.regionCopies[e];
      const int          xp = rc.c[0][0] * N + rc.c[0][1] * x + rc.c[0][2] * y + rc.c[0][3] * z;
      const int          yp = rc.c[1][0] * N + rc.c[1][1] * x + rc.c[1][2] * y + rc.c[1][3] * z;
      const int          zp = rc.c[2][0] * N + rc.c[2][1] * x + rc.c[2][2] * y + rc.c[2][3] * z;
      const double* __restrict__ w = W + ( (size_t) rc.slot * HB_NMASK + rc.maskp ) * HB_NDIR;
      // owned cells: pointer arithmetic (one dependent load less); mirrored cells of other ranks: pointer table
      const double* __restrict__ s = rc.slot < lv.nCells ? srcBase + (size_t) rc.slot * lv.cellSize : src[rc.slot];
      const int                  vm = hb_valid[rc.maskp];
      // linear index of the image point and the index deltas of the six (dy,dz) row changes of the 15-point stencil
      const int ws   = N + 1 - zp;                // width of slice zp
      const int i0   = hb_index( N, xp, yp, zp );
      const int szUp = ( ws * ( ws + 1 ) ) / 2;   // size of slice zp   (distance to slice zp+1)
      const int szDn = ( ( ws + 1 ) * ( ws + 2 ) ) / 2; // size of slice zp-1
      const int dN = ws - yp, dS = -( ws - yp + 1 );        // (0,+1,0), (0,-1,0)
      const int dT = szUp - yp, dTS = szUp - ws;            // (0,0,+1), (0,-1,+1)
      const int dB = yp - szDn, dBN = ws + 1 - szDn;        // (0,0,-1), (0,+1,-1)
      // all neighbour loads are issued together; a direction that leaves the macro-cell has weight 0 in W[.][maskp]
      // (stencil_assembly.cpp)
      int idx[HB_NDIR];
      idx[D_W]   = i0 - 1;
      idx[D_E]   = i0 + 1;
      idx[D_N]   = i0 + dN;
      idx[D_S]   = i0 + dS;
      idx[D_NW]  = i0 + dN - 1;
      idx[D_SE]  = i0 + dS + 1;
      idx[D_TC]  = i0 + dT;
      idx[D_TW]  = i0 + dT - 1;
      idx[D_TS]  = i0 + dTS;
      idx[D_TSE] = i0 + dTS + 1;
      idx[D_BC]  = i0 + dB;
      idx[D_BN]  = i0 + dBN;
      idx[D_BE]  = i0 + dB + 1;
      idx[D_BNW] = i0 + dBN - 1;
      // vm is block-uniform (all threads of a block work on the same region): directions that leave the macro-cell are
      // skipped without divergence, which saves their L1 wavefronts (a face point keeps 9 of its 14 neighbours in a cell)
      double v[HB_NDIR];
#pragma unroll
      for ( int d = 1; d < HB_NDIR; d++ )
         v[d] = ( ( vm >> d ) & 1 ) ? s[idx[d]] : 0.0;
      double part = 0.0;
#pragma unroll
      for ( int d = 1; d < HB_NDIR; d++ )
         part += w[d] * v[d];
      sNb += part;
      diag += w[0];
   }
}

__device__ __forceinline__ void hb_boundary_stencil( const HbLevelView&   lv,
                                                     const double*        W,
                                                     const double* const* src,
                                                     const double*        srcBase,
                                                     const HbCellInfo*    ci,
                                                     int                  mask,
                                                     int                  x,
                                                     int                  y,
                                                     int                  z,
                                                     double&              sNb,
                                                     double&              diag )
{
   hb_boundary_stencil_range( lv, W, src, srcBase, ci->rcBegin[mask], ci->rcBegin[mask + 1], x, y, z, sNb, diag );
}

// SOR update from the full row sum Ax (centre term included), written with explicit fma so that every kernel that uses it rounds
// the same way:  (1 - omega) c + omega ( invd ( rhs - Ax ) + c )
__host__ __device__ __forceinline__ double hb_sor_value( double omega, double Ax, double invd, double centre, double rhs )
{
   return fma( omega, fma( invd, rhs - Ax, centre ), ( 1.0 - omega ) * centre );
}

// Colour of an interface point inside its macro primitive, identical in every macro-cell that holds
// a copy: the barycentric weights are taken w.r.t. the primitive's macro vertices sorted by global id.
//   macro vertex: 0 ; macro edge: parity of the position along the edge (2 colours);
//   macro face: (b1 + 2 b2) mod 3, a proper 3-colouring of the triangular lattice.
__host__ __device__ __forceinline__ int hb_interface_colour_gv( const int32_t* gv, int N, int x, int y, int z );
__host__ __device__ __forceinline__ int hb_interface_colour( const HbCellInfo& ci, int N, int x, int y, int z )
{
   return hb_interface_colour_gv( ci.gv, N, x, y, z );
}
__host__ __device__ __forceinline__ int hb_interface_colour_gv( const int32_t* gv, int N, int x, int y, int z )
{
   const int b[4] = { N - x - y - z, x, y, z };
   int       idx[4];
   int       n = 0;
#pragma unroll
   for ( int j = 0; j < 4; j++ )
      if ( b[j] != 0 )
      {
         // insertion sort by global vertex id
         int k = n++;
         while ( k > 0 && gv[idx[k - 1]] > gv[j] )
         {
            idx[k] = idx[k - 1];
            k--;
         }
         idx[k] = j;
      }
   if ( n == 2 )
      return b[idx[1]] & 1;
   if ( n == 3 )
      return ( b[idx[1]] + 2 * b[idx[2]] ) % 3;
   return 0;
}
