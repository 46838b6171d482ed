// Host-side assembly of the per-macro-cell region stencils W[mask][dir] of the P1 Laplacian.
//
// For a micro-vertex p of a macro-cell with boundary mask m, W[m][d] is the sum, over all micro
// tetrahedra that contain p AND lie inside the macro-cell, of the local stiffness-matrix entry
// coupling p and p+dir[d].  This is HyTeG's calculateStencilInMacroCellForm_new
// (reference src/hyteg/p1functionspace/P1Elements.hpp:465-515) with getNeighboringElements
// (:215-302); the micro-cell types and their vertex offsets are those of
// celldof::macrocell::getMicroVerticesFromMicroCell (reference src/hyteg/volumedofspace/CellDoFIndexing.hpp:155-198).
// The local stiffness matrix is the standard P1 one, K_ij = |T| grad(phi_i).grad(phi_j)
// (reference forms/form_fenics_generated/p1_tet_diffusion.h, used through P1FenicsForm::integrateRow).
#include "stencil_assembly.hpp"

#include <cmath>
#include <stdexcept>

namespace hb {

const int kDir[HB_NDIR][3] = { { 0, 0, 0 },  { -1, 0, 0 }, { 1, 0, 0 },  { 0, 1, 0 },  { 0, -1, 0 },
                               { -1, 1, 0 }, { 1, -1, 0 }, { 0, 0, 1 },  { -1, 0, 1 }, { 0, -1, 1 },
                               { 1, -1, 1 }, { 0, 0, -1 }, { 0, 1, -1 }, { 1, 0, -1 }, { -1, 1, -1 } };

// WHITE_UP, BLUE_UP, GREEN_UP, WHITE_DOWN, BLUE_DOWN, GREEN_DOWN
const int kMicroCell[6][4][3] = { { { 0, 0, 0 }, { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } },
                                  { { 1, 0, 0 }, { 1, 1, 0 }, { 0, 1, 0 }, { 1, 0, 1 } },
                                  { { 1, 0, 0 }, { 0, 1, 0 }, { 1, 0, 1 }, { 0, 0, 1 } },
                                  { { 1, 1, 0 }, { 1, 1, 1 }, { 0, 1, 1 }, { 1, 0, 1 } },
                                  { { 1, 0, 1 }, { 0, 1, 1 }, { 0, 0, 1 }, { 0, 1, 0 } },
                                  { { 0, 1, 0 }, { 1, 1, 0 }, { 1, 0, 1 }, { 0, 1, 1 } } };

int dirIndex( int dx, int dy, int dz )
{
   for ( int d = 0; d < HB_NDIR; d++ )
      if ( kDir[d][0] == dx && kDir[d][1] == dy && kDir[d][2] == dz )
         return d;
   return -1;
}

bool dirValid( int mask, int dx, int dy, int dz )
{
   return !( ( ( mask & 1 ) && dx < 0 ) || ( ( mask & 2 ) && dy < 0 ) || ( ( mask & 4 ) && dz < 0 ) ||
             ( ( mask & 8 ) && dx + dy + dz > 0 ) );
}

unsigned makeValidMask( int mask )
{
   unsigned v = 0;
   for ( int d = 0; d < HB_NDIR; d++ )
      if ( dirValid( mask, kDir[d][0], kDir[d][1], kDir[d][2] ) )
         v |= 1u << d;
   return v;
}

void p1LaplaceElementMatrix( const double p[4][3], double K[4][4] )
{
   // gradients of the barycentric coordinates via the inverse Jacobian
   double J[3][3];
   for ( int i = 0; i < 3; i++ )
      for ( int k = 0; k < 3; k++ )
         J[k][i] = p[i + 1][k] - p[0][k]; // columns = edge vectors
   const double det = J[0][0] * ( J[1][1] * J[2][2] - J[1][2] * J[2][1] ) - J[0][1] * ( J[1][0] * J[2][2] - J[1][2] * J[2][0] ) +
                      J[0][2] * ( J[1][0] * J[2][1] - J[1][1] * J[2][0] );
   if ( det == 0.0 )
      throw std::runtime_error( "hyteg_b200: degenerate tetrahedron" );
   double inv[3][3]; // inverse of J
   inv[0][0] = ( J[1][1] * J[2][2] - J[1][2] * J[2][1] ) / det;
   inv[0][1] = ( J[0][2] * J[2][1] - J[0][1] * J[2][2] ) / det;
   inv[0][2] = ( J[0][1] * J[1][2] - J[0][2] * J[1][1] ) / det;
   inv[1][0] = ( J[1][2] * J[2][0] - J[1][0] * J[2][2] ) / det;
   inv[1][1] = ( J[0][0] * J[2][2] - J[0][2] * J[2][0] ) / det;
   inv[1][2] = ( J[0][2] * J[1][0] - J[0][0] * J[1][2] ) / det;
   inv[2][0] = ( J[1][0] * J[2][1] - J[1][1] * J[2][0] ) / det;
   inv[2][1] = ( J[0][1] * J[2][0] - J[0][0] * J[2][1] ) / det;
   inv[2][2] = ( J[0][0] * J[1][1] - J[0][1] * J[1][0] ) / det;
   // grad(lambda_{i+1}) = row i of inv ; grad(lambda_0) = -sum
   double g[4][3];
   for ( int k = 0; k < 3; k++ )
   {
      g[0][k] = 0.0;
      for ( int i = 0; i < 3; i++ )
      {
         g[i + 1][k] = inv[i][k];
         g[0][k] -= inv[i][k];
      }
   }
   const double vol = std::fabs( det ) / 6.0;
   for ( int i = 0; i < 4; i++ )
      for ( int j = 0; j < 4; j++ )
         K[i][j] = vol * ( g[i][0] * g[j][0] + g[i][1] * g[j][1] + g[i][2] * g[j][2] );
}

void assembleRegionStencils( const double v[4][3], int level, double* W )
{
   const double h = 1.0 / (double) ( 1 << level );
   for ( int i = 0; i < HB_NMASK * HB_NDIR; i++ )
      W[i] = 0.0;
   for ( int t = 0; t < 6; t++ )
   {
      // geometry of a micro-cell of type t anchored at the macro vertex v0 (affine => translation invariant)
      double p[4][3];
      for ( int k = 0; k < 4; k++ )
         for ( int c = 0; c < 3; c++ )
            p[k][c] = v[0][c] + h * ( kMicroCell[t][k][0] * ( v[1][c] - v[0][c] ) + kMicroCell[t][k][1] * ( v[2][c] - v[0][c] ) +
                                      kMicroCell[t][k][2] * ( v[3][c] - v[0][c] ) );
      double K[4][4];
      p1LaplaceElementMatrix( p, K );
      for ( int lv = 0; lv < 4; lv++ ) // the reference micro-vertex is local vertex lv of this element
         for ( int m = 0; m < HB_NMASK; m++ )
         {
            bool inside = true;
            int  dIdx[4];
            for ( int k = 0; k < 4; k++ )
            {
               const int dx = kMicroCell[t][k][0] - kMicroCell[t][lv][0];
               const int dy = kMicroCell[t][k][1] - kMicroCell[t][lv][1];
               const int dz = kMicroCell[t][k][2] - kMicroCell[t][lv][2];
               dIdx[k]      = dirIndex( dx, dy, dz );
               if ( dIdx[k] < 0 )
                  throw std::runtime_error( "hyteg_b200: micro-cell edge outside the 15-point stencil" );
               inside = inside && dirValid( m, dx, dy, dz );
            }
            if ( !inside )
               continue;
            for ( int k = 0; k < 4; k++ )
               W[m * HB_NDIR + dIdx[k]] += K[lv][k];
         }
   }
}

} // namespace hb
