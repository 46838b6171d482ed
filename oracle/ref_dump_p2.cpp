// TEST INFRASTRUCTURE ONLY: P2 / elementwise / blending commands of the reference driver (see oracle/ref_dump.cpp).
//
//   p2apply    P2ElementwiseOperator< Form >::apply on a P2Function (vertexdof + edgedof) handed over as ONE point cloud (micro-vertex
//              positions and micro-edge midpoints, in PHYSICAL coordinates when the shell map is on: P2Function::interpolate evaluates
//              its expression at the blended position); the result is dumped as two clouds (vertex DoFs / edge DoFs) in COMPUTATIONAL
//              coordinates.  form = divkgrad (forms::p2_div_k_grad_blending_q4 with the coefficient k() below), laplace
//              (P2ElementwiseBlendingLaplaceOperator), epsilon (forms::p2_epsilonvar_{i}_{j}_blending_q2, mu = k()).
//   p2restrict / p2prolongate   P2toP2QuadraticRestriction / Prolongation.
//   shellmap   IcosahedralShellMap of every macro-cell: evalF and evalDF at a fixed set of barycentric points.
#include "ref_dump_p2.hpp"

#include <cstdio>
#include <fstream>

#include "hyteg/edgedofspace/EdgeDoFIndexing.hpp"
#include "hyteg/edgedofspace/EdgeDoFMacroCell.hpp"
#include "hyteg/elementwiseoperators/P1ToP2ElementwiseOperator.hpp"
#include "hyteg/elementwiseoperators/P2ElementwiseOperator.hpp"
#include "hyteg/elementwiseoperators/P2ToP1ElementwiseOperator.hpp"
#include "hyteg/forms/form_hyteg_generated/p2/p2_div_k_grad_blending_q4.hpp"
#include "hyteg/forms/form_hyteg_generated/p2/p2_epsilonvar_blending_q2.hpp"
#include "hyteg/geometry/IcosahedralShellMap.hpp"
#include "hyteg/gridtransferoperators/P2toP2QuadraticProlongation.hpp"
#include "hyteg/gridtransferoperators/P2toP2QuadraticRestriction.hpp"
#include "hyteg/p1functionspace/VertexDoFMacroCell.hpp"
#include "hyteg/p2functionspace/P2Function.hpp"
#include "hyteg/solvers/CGSolver.hpp"
#include "hyteg/solvers/ChebyshevSmoother.hpp"
#include "hyteg/solvers/GeometricMultigridSolver.hpp"

using namespace hyteg;
using walberla::real_t;
using walberla::uint_t;

namespace refdump {

void setShellMap( SetupPrimitiveStorage& setup ) { IcosahedralShellMap::setMap( setup ); }

// the variable coefficient of the fixtures (restated in oracle/p2_oracle.py: coefficient_k)
static real_t coefficientK( const Point3D& p )
{
   return 2.0 + std::sin( 3.0 * p[0] ) * std::cos( 2.0 * p[1] ) + p[2] * p[2];
}

static void setP2FromCloud( const P2Function< real_t >& f, const std::string& path, double h, uint_t level )
{
   Cloud                                      c( path, h );
   std::function< real_t( const Point3D& ) > fn = [&c]( const Point3D& p ) { return c( p ); };
   f.interpolate( fn, level, All );
}

static void dumpP2( const P2Function< real_t >& f, uint_t level, const std::string& pathV, const std::string& pathE )
{
   // same order as P2ElementwiseOperator::gemv (P2ElementwiseOperator.cpp:124-128)
   f.communicate< Face, Cell >( level );
   f.communicate< Edge, Cell >( level );
   f.communicate< Vertex, Cell >( level );
   std::vector< double > outV, outE;
   for ( const auto& it : f.getStorage()->getCells() )
   {
      const Cell&   cell  = *it.second;
      const real_t* vdata = cell.getData( f.getVertexDoFFunction().getCellDataID() )->getPointer( level );
      const real_t* edata = cell.getData( f.getEdgeDoFFunction().getCellDataID() )->getPointer( level );
      for ( const auto& idx : vertexdof::macrocell::Iterator( level, 0 ) )
      {
         Point3D p = vertexdof::macrocell::coordinateFromIndex( level, cell, idx );
         outV.insert( outV.end(), { p[0], p[1], p[2], vdata[vertexdof::macrocell::index( level, idx.x(), idx.y(), idx.z() )] } );
      }
      for ( const auto& o : edgedof::allEdgeDoFOrientationsWithoutXYZ )
         for ( const auto& idx : edgedof::macrocell::Iterator( level, 0 ) )
         {
            Point3D p = edgedof::macrocell::coordinateFromIndex( level, cell, idx, o );
            outE.insert( outE.end(), { p[0], p[1], p[2], edata[edgedof::macrocell::index( level, idx.x(), idx.y(), idx.z(), o )] } );
         }
      for ( const auto& idx : edgedof::macrocell::IteratorXYZ( level, 0 ) )
      {
         Point3D p = edgedof::macrocell::coordinateFromIndex( level, cell, idx, edgedof::EdgeDoFOrientation::XYZ );
         outE.insert( outE.end(),
                      { p[0], p[1], p[2], edata[edgedof::macrocell::index( level, idx.x(), idx.y(), idx.z(), edgedof::EdgeDoFOrientation::XYZ )] } );
      }
   }
   std::ofstream ov( pathV, std::ios::binary ), oe( pathE, std::ios::binary );
   ov.write( (const char*) outV.data(), std::streamsize( outV.size() * 8 ) );
   oe.write( (const char*) outE.data(), std::streamsize( outE.size() * 8 ) );
}

template < class OpT >
static void applyAndDump( const OpT& A, const Args& a, const std::shared_ptr< PrimitiveStorage >& st, uint_t L, double h )
{
   P2Function< real_t > src( "src", st, L, L ), dst( "dst", st, L, L );
   setP2FromCloud( src, a.str( "src" ), h, L );
   if ( a.has( "dst" ) )
      setP2FromCloud( dst, a.str( "dst" ), h, L );
   A.apply( src, dst, L, DoFType( int( a.num( "flag", 1 ) ) ), a.num( "update", 0 ) ? Add : Replace );
   dumpP2( dst, L, a.str( "out" ), a.str( "oute" ) );
}

int mainP2( const Args& a )
{
   const uint_t L     = uint_t( a.num( "level", 2 ) );
   const double h     = a.num( "h", 1e-3 );
   const bool   shell = a.num( "shell", 0 ) != 0;
   auto         st    = makeStorage( a, shell );

   if ( a.cmd == "p2apply" )
   {
      const std::string form = a.str( "form", "divkgrad" );
      if ( form == "divkgrad" )
      {
         forms::p2_div_k_grad_blending_q4                          f( coefficientK, coefficientK );
         P2ElementwiseOperator< forms::p2_div_k_grad_blending_q4 > A( st, L, L, f );
         applyAndDump( A, a, st, L, h );
      }
      else if ( form == "laplace" )
      {
         P2ElementwiseBlendingLaplaceOperator A( st, L, L );
         applyAndDump( A, a, st, L, h );
      }
      else if ( form == "epsilon" )
      {
         const int i = int( a.num( "i", 0 ) ), j = int( a.num( "j", 0 ) );
#define EPS_CASE( I, J, ... )                                                                  \
   if ( i == I && j == J )                                                                     \
   {                                                                                           \
      forms::p2_epsilonvar_##I##_##J##_blending_q2                          f( __VA_ARGS__ ); \
      P2ElementwiseOperator< forms::p2_epsilonvar_##I##_##J##_blending_q2 > A( st, L, L, f ); \
      applyAndDump( A, a, st, L, h );                                                          \
   }
         // (the forms that only exist in 3D take one callback, p2_epsilonvar_blending_q2.hpp:53,235)
         EPS_CASE( 0, 0, coefficientK, coefficientK )
         EPS_CASE( 0, 1, coefficientK, coefficientK )
         EPS_CASE( 0, 2, coefficientK )
         EPS_CASE( 1, 0, coefficientK, coefficientK )
         EPS_CASE( 1, 1, coefficientK, coefficientK )
         EPS_CASE( 1, 2, coefficientK )
         EPS_CASE( 2, 0, coefficientK )
         EPS_CASE( 2, 1, coefficientK )
         EPS_CASE( 2, 2, coefficientK )
#undef EPS_CASE
      }
      else
      {
         std::fprintf( stderr, "unknown form %s\n", form.c_str() );
         return 1;
      }
   }
   else if ( a.cmd == "p2top1" || a.cmd == "p2fromp1" )
   {
      // B and B^T blocks of the Stokes operator: P2ToP1ElementwiseBlendingDiv{x,y,z}Operator / P1ToP2ElementwiseBlendingDivT{x,y,z}Operator
      const int            comp = int( a.num( "comp", 0 ) );
      const DoFType        flag = DoFType( int( a.num( "flag", 1 ) ) );
      P2Function< real_t > u( "u", st, L, L );
      P1Function< real_t > q( "q", st, L, L );
      if ( a.cmd == "p2top1" )
      {
         setP2FromCloud( u, a.str( "src" ), h, L );
         if ( comp == 0 )
            P2ToP1ElementwiseBlendingDivxOperator( st, L, L ).apply( u, q, L, flag, Replace );
         else if ( comp == 1 )
            P2ToP1ElementwiseBlendingDivyOperator( st, L, L ).apply( u, q, L, flag, Replace );
         else
            P2ToP1ElementwiseBlendingDivzOperator( st, L, L ).apply( u, q, L, flag, Replace );
         u.getVertexDoFFunction().assign( { 1.0 }, { q }, L, All );
         u.getEdgeDoFFunction().interpolate( real_t( 0 ), L, All );
      }
      else
      {
         Cloud                                      c( a.str( "src" ), h );
         std::function< real_t( const Point3D& ) > fn = [&c]( const Point3D& p ) { return c( p ); };
         q.interpolate( fn, L, All );
         if ( comp == 0 )
            P1ToP2ElementwiseBlendingDivTxOperator( st, L, L ).apply( q, u, L, flag, Replace );
         else if ( comp == 1 )
            P1ToP2ElementwiseBlendingDivTyOperator( st, L, L ).apply( q, u, L, flag, Replace );
         else
            P1ToP2ElementwiseBlendingDivTzOperator( st, L, L ).apply( q, u, L, flag, Replace );
      }
      dumpP2( u, L, a.str( "out" ), a.str( "oute" ) );
   }
   else if ( a.cmd == "p2gmg" || a.cmd == "p2cheb" )
   {
      // TerraNeo-style recipe on the scalar blended div-k-grad operator: GeometricMultigridSolver with ChebyshevSmoother (inverse
      // diagonal, radius from power iterations on the finest level), CG on the coarsest level, quadratic transfer
      using OpT           = P2ElementwiseOperator< forms::p2_div_k_grad_blending_q4 >;
      const uint_t minL   = uint_t( a.num( "minlevel", 1 ) );
      forms::p2_div_k_grad_blending_q4 f( coefficientK, coefficientK );
      OpT                              A( st, minL, L, f );
      A.computeInverseDiagonalOperatorValues();
      P2Function< real_t > x( "x", st, minL, L ), b( "b", st, minL, L ), r( "r", st, minL, L ), ev( "ev", st, minL, L ), tmp( "tmp", st, minL, L ),
          one( "one", st, minL, L );
      setP2FromCloud( x, a.str( "x" ), h, L );
      setP2FromCloud( b, a.str( "rhs" ), h, L );
      setP2FromCloud( ev, a.str( "ev" ), h, L );
      one.interpolate( real_t( 1 ), L, All );
      const double npts = one.dotGlobal( one, L, Inner );
      const double rho  = chebyshev::estimateRadius( A, L, uint_t( a.num( "power_iterations", 20 ) ), st, ev, tmp );
      std::printf( "RESULT rho %.17g\n", rho );
      auto cheb = std::make_shared< ChebyshevSmoother< OpT > >( st, minL, L );
      cheb->setupCoefficients( uint_t( a.num( "order", 3 ) ), rho );
      auto res2 = [&]() {
         A.apply( x, r, L, Inner );
         r.assign( { 1.0, -1.0 }, { b, r }, L, Inner );
         return r.dotGlobal( r, L, Inner ) / npts;
      };
      if ( a.cmd == "p2cheb" )
      {
         for ( int k = 0; k < int( a.num( "sweeps", 1 ) ); k++ )
            cheb->solve( A, x, b, L );
      }
      else
      {
         auto coarse = std::make_shared< CGSolver< OpT > >( st, minL, minL, 1000, 1e-14, 1e-14 );
         auto R      = std::make_shared< P2toP2QuadraticRestriction >();
         auto P      = std::make_shared< P2toP2QuadraticProlongation >();
         GeometricMultigridSolver< OpT > gmg( st, cheb, coarse, R, P, minL, L, uint_t( a.num( "pre", 2 ) ), uint_t( a.num( "post", 2 ) ) );
         std::printf( "RESULT res2 %.17g\n", res2() );
         for ( int c = 0; c < int( a.num( "cycles", 5 ) ); c++ )
         {
            gmg.solve( A, x, b, L );
            std::printf( "RESULT res2 %.17g\n", res2() );
         }
      }
      dumpP2( x, L, a.str( "out" ), a.str( "oute" ) );
   }
   else if ( a.cmd == "p2restrict" || a.cmd == "p2prolongate" )
   {
      // level = source level
      const uint_t         lo = a.cmd == "p2restrict" ? L - 1 : L, hi = lo + 1;
      auto                 st2 = st;
      P2Function< real_t > f( "f", st2, lo, hi );
      setP2FromCloud( f, a.str( "src" ), h, L );
      const DoFType flag = DoFType( int( a.num( "flag", 5 ) ) );
      if ( a.cmd == "p2restrict" )
         P2toP2QuadraticRestriction().restrict( f, L, flag );
      else
         P2toP2QuadraticProlongation().prolongate( f, L, flag );
      dumpP2( f, a.cmd == "p2restrict" ? lo : hi, a.str( "out" ), a.str( "oute" ) );
   }
   else if ( a.cmd == "shellmap" )
   {
      // per macro-cell: 4 vertices (12), then for each of the sample points (barycentric) F (3) and DF (9, row-major)
      static const double bary[5][4] = { { 0.25, 0.25, 0.25, 0.25 }, { 0.7, 0.1, 0.1, 0.1 }, { 0.1, 0.6, 0.2, 0.1 }, { 0.05, 0.15, 0.3, 0.5 }, { 1, 0, 0, 0 } };
      std::vector< double > out;
      for ( const auto& it : st->getCells() )
      {
         const Cell& cell = *it.second;
         for ( int k = 0; k < 4; k++ )
            for ( int d = 0; d < 3; d++ )
               out.push_back( cell.getCoordinates()[uint_t( k )][d] );
         for ( int s = 0; s < 5; s++ )
         {
            Point3D x( 0, 0, 0 ), y;
            for ( int k = 0; k < 4; k++ )
               x += bary[s][k] * cell.getCoordinates()[uint_t( k )];
            Matrix3r DF;
            cell.getGeometryMap()->evalF( x, y );
            cell.getGeometryMap()->evalDF( x, DF );
            out.insert( out.end(), { y[0], y[1], y[2] } );
            for ( int r = 0; r < 3; r++ )
               for ( int c = 0; c < 3; c++ )
                  out.push_back( DF( r, c ) );
         }
      }
      std::ofstream o( a.str( "out" ), std::ios::binary );
      o.write( (const char*) out.data(), std::streamsize( out.size() * 8 ) );
   }
   else
   {
      std::fprintf( stderr, "unknown command %s\n", a.cmd.c_str() );
      return 1;
   }
   return 0;
}

} // namespace refdump
