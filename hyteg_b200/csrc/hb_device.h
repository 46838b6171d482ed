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
   const double*        W;   // [nSlots][16][15]
   double*              dst; // local arena
   const double*        rhs; // local arena
   double*              aux; // local arena
   double*              scratch; // [nCells*nBnd] staging for interface phases (SOR)
   double               omega;
   int32_t              flag;    // DoFType selection
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

// Stencil at a point on the boundary of macro-cell 'cell' (local index): sum of the partial
// stencils of every macro-cell that contains the point, visited in ascending global cell id so
// that all copies of the point produce bit-identical results.  The partial stencil of one cell is
// exactly HyTeG's calculateStencilInMacroCellForm_new (reference P1Elements.hpp:465-515); the sum
// over the neighbouring cells is how HyTeG builds its macro-face/edge/vertex stencils
// (reference P1Elements.hpp:626-720, 870-1017).
__device__ __forceinline__ void hb_boundary_stencil( const HbLevelView&   lv,
                                                     const double*        W,
                                                     const double* const* src,
                                                     int                  cell,
                                                     int                  x,
                                                     int                  y,
                                                     int                  z,
                                                     double&              sNb,
                                                     double&              diag )
{
   const int        N  = lv.N;
   const HbCellInfo ci = lv.cells[cell];
   const int        b[4] = { N - x - y - z, x, y, z };
   int              supp = 0;
#pragma unroll
   for ( int j = 0; j < 4; j++ )
      supp |= ( b[j] != 0 ) << j;
   sNb  = 0.0;
   diag = 0.0;
   for ( int e = ci.copyBegin; e < ci.copyEnd; e++ )
   {
      const HbCopy cp = lv.copies[e];
      if ( supp & ~cp.sharedMask )
         continue;
      int bp[4] = { 0, 0, 0, 0 };
#pragma unroll
      for ( int j = 0; j < 4; j++ )
         if ( ( supp >> j ) & 1 )
            bp[cp.vmap[j]] = b[j];
      const int     xp = bp[1], yp = bp[2], zp = bp[3];
      const int     mp = hb_mask( N, xp, yp, zp );
      const double* w  = W + ( (size_t) cp.slot * HB_NMASK + mp ) * HB_NDIR;
      const double* s  = src[cp.slot];
      const int     vm = hb_valid[mp];
      double        part = 0.0;
#pragma unroll
      for ( int d = 1; d < HB_NDIR; d++ )
         if ( ( vm >> d ) & 1 )
            part += w[d] * s[hb_index( N, xp + hb_dir[d][0], yp + hb_dir[d][1], zp + hb_dir[d][2] )];
      sNb += part;
      diag += w[0];
   }
}

// Colour of an interface point inside its macro primitive, identical in every macro-cell that holds
// a copy: the barycentric weights are taken w.r.t. the primitive's macro vertices sorted by global id.
//   macro vertex: 0 ; macro edge: parity of the position along the edge (2 colours);
//   macro face: (b1 + 2 b2) mod 3, a proper 3-colouring of the triangular lattice.
__device__ __forceinline__ int hb_interface_colour( const HbCellInfo& ci, int N, int x, int y, int z )
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
         while ( k > 0 && ci.gv[idx[k - 1]] > ci.gv[j] )
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
