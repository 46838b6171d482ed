#pragma once
#include "hb_types.h"

namespace hb {

extern const int kDir[HB_NDIR][3];
extern const int kMicroCell[6][4][3];

int      dirIndex( int dx, int dy, int dz );
bool     dirValid( int mask, int dx, int dy, int dz );
unsigned makeValidMask( int mask );

// P1 Laplace element matrix of the tetrahedron p[0..3]
void p1LaplaceElementMatrix( const double p[4][3], double K[4][4] );

// W[16][15] for the macro-cell with vertex coordinates v[0..3] on refinement level 'level'
void assembleRegionStencils( const double v[4][3], int level, double* W );

} // namespace hb
