// C entry points around the REFERENCE's own generated macro-cell kernels, compiled from the sources
// where they lie under /root/reference (see oracle/Makefile).  Test infrastructure / CPU baseline only.
#include <map>

#include "constant_stencil_operator/P1generatedKernels/apply_3D_macrocell_vertexdof_to_vertexdof_add.hpp"
#include "constant_stencil_operator/P1generatedKernels/apply_3D_macrocell_vertexdof_to_vertexdof_replace.hpp"
#include "constant_stencil_operator/P1generatedKernels/sor_3D_macrocell_P1.hpp"
#include "constant_stencil_operator/P1generatedKernels/sor_3D_macrocell_P1_backwards.hpp"
#include "hyteg/gridtransferoperators/generatedKernels/prolongate_3D_macrocell_P1_push_additive.hpp"
#include "hyteg/gridtransferoperators/generatedKernels/restrict_3D_macrocell_P1_pull_additive.hpp"

using hyteg::indexing::Index;

// direction order of hyteg_b200: C W E N S NW SE TC TW TS TSE BC BN BE BNW
static const int kDir[15][3] = { { 0, 0, 0 },  { -1, 0, 0 }, { 1, 0, 0 },  { 0, 1, 0 },  { 0, -1, 0 },
                                 { -1, 1, 0 }, { 1, -1, 0 }, { 0, 0, 1 },  { -1, 0, 1 }, { 0, -1, 1 },
                                 { 1, -1, 1 }, { 0, 0, -1 }, { 0, 1, -1 }, { 1, 0, -1 }, { -1, 1, -1 } };

static std::map< Index, double > stencilMap( const double* w15 )
{
   std::map< Index, double > m;
   for ( int d = 0; d < 15; d++ )
      m[Index( kDir[d][0], kDir[d][1], kDir[d][2] )] = w15[d];
   return m;
}

extern "C" {

void ref_apply_macrocell( double* dst, const double* src, int level, const double* w15, int add )
{
   auto m = stencilMap( w15 );
   if ( add )
      hyteg::vertexdof::macrocell::generated::apply_3D_macrocell_vertexdof_to_vertexdof_add( dst, src, level, m );
   else
      hyteg::vertexdof::macrocell::generated::apply_3D_macrocell_vertexdof_to_vertexdof_replace( dst, src, level, m );
}

void ref_sor_macrocell( double* dst, const double* rhs, int level, const double* w15, double relax, int backwards )
{
   auto m = stencilMap( w15 );
   if ( backwards )
      hyteg::vertexdof::macrocell::generated::sor_3D_macrocell_P1_backwards( dst, rhs, level, m, relax );
   else
      hyteg::vertexdof::macrocell::generated::sor_3D_macrocell_P1( dst, rhs, level, m, relax );
}

// nn[14]: numNeighborCells of edges 0..5, faces 0..3, vertices 0..3
void ref_restrict_macrocell( double* coarse, const double* fine, int coarseLevel, const double* nn )
{
   hyteg::vertexdof::macrocell::generated::restrict_3D_macrocell_P1_pull_additive(
       coarse, fine, coarseLevel, nn[0], nn[1], nn[2], nn[3], nn[4], nn[5], nn[6], nn[7], nn[8], nn[9], nn[10], nn[11], nn[12], nn[13] );
}

void ref_prolongate_macrocell( const double* coarse, double* fine, int coarseLevel, const double* nn )
{
   hyteg::vertexdof::macrocell::generated::prolongate_3D_macrocell_P1_push_additive(
       coarse, fine, coarseLevel, nn[0], nn[1], nn[2], nn[3], nn[4], nn[5], nn[6], nn[7], nn[8], nn[9], nn[10], nn[11], nn[12], nn[13] );
}
}
