// TEST INFRASTRUCTURE ONLY.  A small command-line driver written against the REFERENCE's public C++ API (HyTeG), linked against
// oracle/_ref/libhyteg_ref.a (= the reference's own sources compiled by oracle/build_ref.py).  It runs the reference's
// P1ConstantLaplaceOperator / smoothers / transfer operators / GeometricMultigridSolver on data handed over as point clouds and
// writes the results back as point clouds, so that oracle/make_ref_fixtures.py can turn them into tests/golden/ref_*.npz and
// bench.py --impl reference can time the reference's real apply.  Nothing of the product links or calls this.
//
// Data exchange: a "cloud" file is n records of 4 doubles (x, y, z, value).  Input functions are set with the reference's own
// VertexDoFFunction::interpolate( std::function ) which looks the value up by coordinate; outputs are dumped from the macro-cell
// arrays (closed macro-tet, after communicating the lower-dimensional primitives' values up to the cells).  No assumption about
// the reference's internal primitive numbering or orientation enters.
//
// Usage: ref_dump <command> key=value ...   (see main)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "core/Environment.h"
#include "core/mpi/MPIManager.h"

#include "hyteg/gridtransferoperators/P1toP1LinearProlongation.hpp"
#include "hyteg/gridtransferoperators/P1toP1LinearRestriction.hpp"
#include "hyteg/mesh/MeshInfo.hpp"
#include "hyteg/p1functionspace/P1Function.hpp"
#include "hyteg/p1functionspace/VertexDoFMacroCell.hpp"
#include "hyteg/primitivestorage/PrimitiveStorage.hpp"
#include "hyteg/primitivestorage/SetupPrimitiveStorage.hpp"
#include "hyteg/solvers/CGSolver.hpp"
#include "hyteg/solvers/ChebyshevSmoother.hpp"
#include "hyteg/solvers/GaussSeidelSmoother.hpp"
#include "hyteg/solvers/GeometricMultigridSolver.hpp"
#include "hyteg/solvers/SORSmoother.hpp"
#include "hyteg/solvers/SymmetricGaussSeidelSmoother.hpp"
#include "hyteg/solvers/WeightedJacobiSmoother.hpp"

#include "constant_stencil_operator/P1ConstantOperator.hpp"

#include "ref_dump_p2.hpp"

using namespace hyteg;
using walberla::real_t;
using walberla::uint_t;

namespace refdump {

double Args::num( const std::string& k, double dflt ) const
{
   auto it = kv.find( k );
   return it == kv.end() ? dflt : std::atof( it->second.c_str() );
}
std::string Args::str( const std::string& k, const std::string& dflt ) const
{
   auto it = kv.find( k );
   return it == kv.end() ? dflt : it->second;
}
bool Args::has( const std::string& k ) const { return kv.count( k ) != 0; }

Cloud::Cloud( const std::string& path, double bucket )
: h( bucket )
{
   std::ifstream f( path, std::ios::binary );
   if ( !f )
   {
      std::fprintf( stderr, "cannot open %s\n", path.c_str() );
      std::exit( 2 );
   }
   f.seekg( 0, std::ios::end );
   size_t bytes = size_t( f.tellg() );
   f.seekg( 0 );
   d.resize( bytes / 8 );
   f.read( (char*) d.data(), std::streamsize( bytes ) );
   for ( size_t i = 0; i < d.size() / 4; i++ )
      buckets[key( d[4 * i], d[4 * i + 1], d[4 * i + 2] )].push_back( int( i ) );
}
uint64_t Cloud::key( double x, double y, double z ) const
{
   auto q = [this]( double v ) { return uint64_t( int64_t( std::floor( v / h ) ) + ( 1 << 20 ) ) & 0x1fffff; };
   return ( q( x ) << 42 ) | ( q( y ) << 21 ) | q( z );
}
double Cloud::operator()( const Point3D& p ) const
{
   double best = 1e300, val = 0;
   for ( int dx = -1; dx <= 1; dx++ )
      for ( int dy = -1; dy <= 1; dy++ )
         for ( int dz = -1; dz <= 1; dz++ )
         {
            auto it = buckets.find( key( p[0] + dx * h, p[1] + dy * h, p[2] + dz * h ) );
            if ( it == buckets.end() )
               continue;
            for ( int i : it->second )
            {
               double a = d[4 * i] - p[0], b = d[4 * i + 1] - p[1], c = d[4 * i + 2] - p[2];
               double r = a * a + b * b + c * c;
               if ( r < best )
               {
                  best = r;
                  val  = d[4 * i + 3];
               }
            }
         }
   if ( best > 0.01 * h * h )
   {
      std::fprintf( stderr, "cloud lookup failed at (%g,%g,%g): nearest %g, bucket %g\n", p[0], p[1], p[2], std::sqrt( best ), h );
      std::exit( 3 );
   }
   return val;
}

std::shared_ptr< PrimitiveStorage > makeStorage( const Args& a, bool blending )
{
   auto                  meshInfo = MeshInfo::fromGmshFile( a.str( "mesh" ) );
   SetupPrimitiveStorage setup( meshInfo, 1 );
   if ( blending )
      setShellMap( setup );
   setup.setMeshBoundaryFlagsOnBoundary( 1, 0, true );
   return std::make_shared< PrimitiveStorage >( setup );
}

void setFromCloud( const P1Function< real_t >& f, const std::string& path, double h, uint_t level )
{
   Cloud                                      c( path, h );
   std::function< real_t( const Point3D& ) > fn = [&c]( const Point3D& p ) { return c( p ); };
   f.interpolate( fn, level, All );
}

void dumpP1( const P1Function< real_t >& f, uint_t level, const std::string& path )
{
   f.communicate< Vertex, Edge >( level );
   f.communicate< Edge, Face >( level );
   f.communicate< Face, Cell >( level );
   std::vector< double > out;
   for ( const auto& it : f.getStorage()->getCells() )
   {
      const Cell&   cell = *it.second;
      const real_t* data = cell.getData( f.getCellDataID() )->getPointer( level );
      for ( const auto& idx : vertexdof::macrocell::Iterator( level, 0 ) )
      {
         Point3D p = vertexdof::macrocell::coordinateFromIndex( level, cell, idx );
         out.push_back( p[0] );
         out.push_back( p[1] );
         out.push_back( p[2] );
         out.push_back( data[vertexdof::macrocell::index( level, idx.x(), idx.y(), idx.z() )] );
      }
   }
   std::ofstream o( path, std::ios::binary );
   o.write( (const char*) out.data(), std::streamsize( out.size() * 8 ) );
}

} // namespace refdump

using namespace refdump;
using Op = P1ConstantLaplaceOperator;

// ChebyshevSmoother keeps its coefficients protected: a derived class reads them out
struct ChebPeek : ChebyshevSmoother< Op >
{
   using ChebyshevSmoother< Op >::ChebyshevSmoother;
   const std::vector< real_t >& coeffs( uint_t i ) const { return coefficients[i]; }
};

static double now()
{
   return std::chrono::duration< double >( std::chrono::steady_clock::now().time_since_epoch() ).count();
}

// squared discrete residual of the reference's convergence tests: (A u - f restricted to Inner)^2 / #inner points
// (tests/hyteg/convergence/P1GMG3DConvergenceTest.cpp:121-136 with f = 0; here general f)
static double residual2( const Op& A, const P1Function< real_t >& u, const P1Function< real_t >& f, P1Function< real_t >& r, uint_t L,
                         double npts )
{
   A.apply( u, r, L, Inner );
   r.assign( { 1.0, -1.0 }, { f, r }, L, Inner );
   return r.dotGlobal( r, L, Inner ) / npts;
}

int main( int argc, char** argv )
{
   int                   argcEnv = 1; // the remaining arguments are this driver's, not a waLBerla parameter file
   walberla::Environment env( argcEnv, argv );
   walberla::MPIManager::instance()->useWorldComm();
   walberla::logging::Logging::instance()->setLogLevel( walberla::logging::Logging::WARNING );
   if ( argc < 2 )
   {
      std::fprintf( stderr, "usage: ref_dump <command> key=value ...\n" );
      return 1;
   }
   Args a;
   a.cmd = argv[1];
   for ( int i = 2; i < argc; i++ )
   {
      std::string s( argv[i] );
      auto        p = s.find( '=' );
      if ( p != std::string::npos )
         a.kv[s.substr( 0, p )] = s.substr( p + 1 );
   }
   if ( a.cmd.rfind( "p2", 0 ) == 0 || a.cmd == "shellmap" )
      return mainP2( a );

   const uint_t L    = uint_t( a.num( "level", 3 ) );
   const uint_t minL = uint_t( a.num( "minlevel", double( L ) ) );
   const double h    = a.num( "h", 1e-3 );
   auto         st   = makeStorage( a, false );
   const DoFType flag = DoFType( int( a.num( "flag", 1 ) ) );

   if ( a.cmd == "apply" )
   {
      Op                   A( st, L, L );
      P1Function< real_t > src( "src", st, L, L ), dst( "dst", st, L, L );
      setFromCloud( src, a.str( "src" ), h, L );
      if ( a.has( "dst" ) )
         setFromCloud( dst, a.str( "dst" ), h, L );
      A.apply( src, dst, L, flag, a.num( "update", 0 ) ? Add : Replace );
      dumpP1( dst, L, a.str( "out" ) );
   }
   else if ( a.cmd == "sor" || a.cmd == "jac" || a.cmd == "cheb" || a.cmd == "sgs" )
   {
      Op                   A( st, L, L );
      A.computeInverseDiagonalOperatorValues();
      P1Function< real_t > x( "x", st, L, L ), b( "b", st, L, L ), tmp( "tmp", st, L, L ), ev( "ev", st, L, L );
      setFromCloud( x, a.str( "x" ), h, L );
      setFromCloud( b, a.str( "rhs" ), h, L );
      const int sweeps = int( a.num( "sweeps", 1 ) );
      if ( a.cmd == "sor" )
      {
         for ( int s = 0; s < sweeps; s++ )
            if ( a.num( "backwards", 0 ) )
               A.smooth_sor_backwards( x, b, a.num( "omega", 1.0 ), L, flag );
            else
               A.smooth_sor( x, b, a.num( "omega", 1.0 ), L, flag );
      }
      else if ( a.cmd == "sgs" )
      {
         SymmetricGaussSeidelSmoother< Op > s;
         for ( int k = 0; k < sweeps; k++ )
            s.solve( A, x, b, L );
      }
      else if ( a.cmd == "jac" )
      {
         WeightedJacobiSmoother< Op > s( st, L, L, a.num( "omega", 2.0 / 3.0 ) );
         for ( int k = 0; k < sweeps; k++ )
            s.solve( A, x, b, L );
      }
      else
      {
         setFromCloud( ev, a.str( "ev" ), h, L );
         const double rho = chebyshev::estimateRadius( A, L, uint_t( a.num( "power_iterations", 20 ) ), st, ev, tmp );
         ChebyshevSmoother< Op > s( st, L, L );
         s.setupCoefficients( uint_t( a.num( "order", 3 ) ), rho );
         for ( int k = 0; k < sweeps; k++ )
            s.solve( A, x, b, L );
         std::printf( "RESULT rho %.17g\n", rho );
      }
      dumpP1( x, L, a.str( "out" ) );
   }
   else if ( a.cmd == "cg" )
   {
      Op                   A( st, L, L );
      P1Function< real_t > x( "x", st, L, L ), b( "b", st, L, L ), r( "r", st, L, L ), one( "one", st, L, L );
      setFromCloud( x, a.str( "x" ), h, L );
      setFromCloud( b, a.str( "rhs" ), h, L );
      one.interpolate( 1.0, L, All );
      const double  npts = one.dotGlobal( one, L, Inner );
      CGSolver< Op > cg( st, L, L, uint_t( a.num( "maxiter", 10000 ) ), a.num( "rtol", 1e-16 ), a.num( "atol", 1e-16 ) );
      cg.solve( A, x, b, L );
      std::printf( "RESULT res2 %.17g\n", residual2( A, x, b, r, L, npts ) );
      dumpP1( x, L, a.str( "out" ) );
   }
   else if ( a.cmd == "gmg" )
   {
      Op                   A( st, minL, L );
      A.computeInverseDiagonalOperatorValues();
      P1Function< real_t > x( "x", st, minL, L ), b( "b", st, minL, L ), r( "r", st, minL, L ), one( "one", st, minL, L ),
          ev( "ev", st, minL, L ), tmp( "tmp", st, minL, L );
      setFromCloud( x, a.str( "x" ), h, L );
      setFromCloud( b, a.str( "rhs" ), h, L );
      one.interpolate( 1.0, L, All );
      const double                 npts = one.dotGlobal( one, L, Inner );
      std::shared_ptr< Solver< Op > > sm;
      const std::string            kind = a.str( "smoother", "gs" );
      if ( kind == "gs" )
         sm = std::make_shared< GaussSeidelSmoother< Op > >();
      else if ( kind == "sgs" )
         sm = std::make_shared< SymmetricGaussSeidelSmoother< Op > >();
      else if ( kind == "sor" )
         sm = std::make_shared< SORSmoother< Op > >( a.num( "omega", 1.0 ) );
      else if ( kind == "jac" )
         sm = std::make_shared< WeightedJacobiSmoother< Op > >( st, minL, L, a.num( "omega", 2.0 / 3.0 ) );
      else
      {
         // TerraNeo-style recipe: power iteration on the coarsest smoothed level, same radius on all levels
         auto cheb = std::make_shared< ChebyshevSmoother< Op > >( st, minL, L );
         setFromCloud( ev, a.str( "ev" ), h, L );
         const double rho = chebyshev::estimateRadius( A, L, uint_t( a.num( "power_iterations", 20 ) ), st, ev, tmp );
         cheb->setupCoefficients( uint_t( a.num( "order", 3 ) ), rho );
         std::printf( "RESULT rho %.17g\n", rho );
         sm = cheb;
      }
      auto coarse = std::make_shared< CGSolver< Op > >( st, minL, minL );
      auto R      = std::make_shared< P1toP1LinearRestriction<> >();
      auto P      = std::make_shared< P1toP1LinearProlongation<> >();
      GeometricMultigridSolver< Op > gmg( st, sm, coarse, R, P, minL, L, uint_t( a.num( "pre", 3 ) ), uint_t( a.num( "post", 3 ) ) );
      std::printf( "RESULT res2 %.17g\n", residual2( A, x, b, r, L, npts ) );
      for ( int c = 0; c < int( a.num( "cycles", 4 ) ); c++ )
      {
         gmg.solve( A, x, b, L );
         std::printf( "RESULT res2 %.17g\n", residual2( A, x, b, r, L, npts ) );
      }
      dumpP1( x, L, a.str( "out" ) );
   }
   else if ( a.cmd == "chebcoeffs" )
   {
      // ChebyshevSmoother::setupCoefficients( order, lower, upper ) for orders 1..10 (ChebyshevSmoother.hpp:278-539)
      for ( uint_t order = 1; order <= 10; order++ )
      {
         ChebPeek cheb( st, L, L );
         cheb.setupCoefficientsInternal( order, a.num( "lower", 0.3 ), a.num( "upper", 1.2 ), L );
         for ( real_t c : cheb.coeffs( 0 ) )
            std::printf( "RESULT order%d %.17g\n", int( order ), c );
      }
   }
   else if ( a.cmd == "radius" )
   {
      Op                   A( st, L, L );
      A.computeInverseDiagonalOperatorValues();
      P1Function< real_t > x( "x", st, L, L ), tmp( "tmp", st, L, L );
      setFromCloud( x, a.str( "x" ), h, L );
      std::printf( "RESULT rho %.17g\n", chebyshev::estimateRadius( A, L, uint_t( a.num( "iterations", 20 ) ), st, x, tmp ) );
   }
   else if ( a.cmd == "restrict" || a.cmd == "prolongate" )
   {
      // level = source level
      const uint_t         lo = a.cmd == "restrict" ? L - 1 : L, hi = lo + 1;
      P1Function< real_t > f( "f", st, lo, hi );
      setFromCloud( f, a.str( "src" ), h, L );
      if ( a.cmd == "restrict" )
         P1toP1LinearRestriction<>().restrict( f, L, flag );
      else if ( a.num( "add", 0 ) )
      {
         setFromCloud( f, a.str( "dst" ), h, hi );
         P1toP1LinearProlongation<>().prolongateAndAdd( f, L, flag );
      }
      else
         P1toP1LinearProlongation<>().prolongate( f, L, flag );
      dumpP1( f, a.cmd == "restrict" ? lo : hi, a.str( "out" ) );
   }
   else if ( a.cmd == "bench" )
   {
      // the reference's stock timed call (apps/benchmarks/ApplyBenchmark/ApplyBenchmark.cpp:87): A.apply( src, dst, level, Inner )
      Op                   A( st, minL, L );
      A.computeInverseDiagonalOperatorValues();
      P1Function< real_t > src( "src", st, minL, L ), dst( "dst", st, minL, L ), b( "b", st, minL, L );
      std::function< real_t( const Point3D& ) > fn = []( const Point3D& p ) { return std::sin( p[0] ) * std::sinh( p[1] ) + p[2]; };
      src.interpolate( fn, L, All );
      const std::string what  = a.str( "what", "apply" );
      const int         iters = int( a.num( "iters", 5 ) );
      const double      budget = a.num( "seconds", 1e30 );
      auto              sm    = std::make_shared< GaussSeidelSmoother< Op > >();
      auto              coarse = std::make_shared< CGSolver< Op > >( st, minL, minL );
      auto              R     = std::make_shared< P1toP1LinearRestriction<> >();
      auto              P     = std::make_shared< P1toP1LinearProlongation<> >();
      GeometricMultigridSolver< Op > gmg( st, sm, coarse, R, P, minL, L, 3, 3 );
      auto one = [&]() {
         if ( what == "apply" )
            A.apply( src, dst, L, Inner );
         else if ( what == "sor" )
            A.smooth_sor( src, b, 1.0, L, Inner );
         else if ( what == "jac" )
            A.smooth_jac( dst, b, src, 2.0 / 3.0, L, Inner );
         else
            gmg.solve( A, src, b, L );
      };
      one(); // warm-up
      double t0 = now(), t = t0;
      int    n  = 0;
      while ( n < iters && t - t0 < budget )
      {
         one();
         n++;
         t = now();
      }
      std::printf( "RESULT seconds_per_call %.9g\nRESULT calls %d\nRESULT dofs %.17g\n", ( t - t0 ) / n, n,
                   double( numberOfGlobalDoFs< P1FunctionTag >( *st, L ) ) );
   }
   else
   {
      std::fprintf( stderr, "unknown command %s\n", a.cmd.c_str() );
      return 1;
   }
   return 0;
}
