// C++ host program against include/hyteg_b200/hyteg.hpp, written the way a HyTeG application is written
// (cf. the reference's tests/hyteg/P1 GMG convergence test and P1 Laplace operator test): same class names, same call
// signatures, device library underneath.  Built and run by tests/test_gpu_cpp_dropin.py on the GPU box:
//   g++ -std=c++17 -Iinclude tests/cpp/p1_gmg_dropin.cpp -Lhyteg_b200 -lhyteg_b200 -o p1_gmg_dropin
// Exit code 0 = all checks passed.
#include <cmath>
#include <cstdio>
#include <functional>
#include <memory>

#include "hyteg_b200/hyteg.hpp"

using namespace hyteg;

#define CHECK( cond, msg )                                        \
   do                                                             \
   {                                                              \
      if ( !( cond ) )                                            \
      {                                                           \
         std::fprintf( stderr, "CHECK FAILED: %s\n", msg );       \
         return 1;                                                \
      }                                                           \
   } while ( 0 )

int main( int argc, char** argv )
{
   std::setvbuf( stdout, nullptr, _IOLBF, 0 );
   const uint_t minLevel = 0;
   const uint_t maxLevel = argc > 1 ? (uint_t) std::atoi( argv[1] ) : 5;
   if ( hb_device_init( 0 ) != 0 )
   {
      std::fprintf( stderr, "%s\n", hb_last_error() );
      return 2;
   }

   const auto            meshInfo = MeshInfo::meshCube6el();
   SetupPrimitiveStorage setupStorage( meshInfo, 1 );
   setupStorage.setMeshBoundaryFlagsOnBoundary( 1, 0, true );
   const auto storage = std::make_shared< PrimitiveStorage >( setupStorage );
   CHECK( storage->hasGlobalCells(), "3D storage" );

   P1ConstantLaplaceOperator L( storage, minLevel, maxLevel );

   std::function< real_t( const Point3D& ) > exact  = []( const Point3D& p ) { return std::sin( p[0] ) * std::sinh( p[1] ) + p[2]; };
   std::function< real_t( const Point3D& ) > linear = []( const Point3D& p ) { return 2 * p[0] + 3 * p[1] - 7 * p[2] + 1; };

   P1Function< real_t > u( "u", storage, minLevel, maxLevel ), f( "f", storage, minLevel, maxLevel ), r( "r", storage, minLevel, maxLevel ),
       uExact( "uExact", storage, minLevel, maxLevel ), err( "err", storage, minLevel, maxLevel ), one( "one", storage, minLevel, maxLevel );

   one.interpolate( real_t( 1 ), maxLevel, All );
   const real_t nInner = one.dotGlobal( one, maxLevel, Inner );

   // the operator annihilates linear functions on inner DoFs
   u.interpolate( linear, maxLevel, All );
   L.apply( u, r, maxLevel, Inner );
   const real_t resLinear = std::sqrt( r.dotGlobal( r, maxLevel, Inner ) / nInner );
   std::printf( "A * linear: discrete L2 = %.3e\n", resLinear );
   CHECK( resLinear < 2.8e-13, "A * linear != 0" );

   // V(3,3) cycles, Gauss-Seidel smoother, CG on the coarsest level
   u.interpolate( real_t( 0 ), maxLevel, Inner );
   u.interpolate( exact, maxLevel, DirichletBoundary );
   f.interpolate( real_t( 0 ), maxLevel, All );
   uExact.interpolate( exact, maxLevel, All );

   auto smoother     = std::make_shared< GaussSeidelSmoother< P1ConstantLaplaceOperator > >();
   auto coarseSolver = std::make_shared< CGSolver< P1ConstantLaplaceOperator > >( storage, minLevel, minLevel );
   auto restriction  = std::make_shared< P1toP1LinearRestriction<> >();
   auto prolongation = std::make_shared< P1toP1LinearProlongation<> >();
   GeometricMultigridSolver< P1ConstantLaplaceOperator > gmg( storage, smoother, coarseSolver, restriction, prolongation, minLevel, maxLevel, 3, 3 );

   L.apply( u, r, maxLevel, Inner );
   real_t last = r.dotGlobal( r, maxLevel, Inner ) / nInner;
   for ( int cycle = 1; cycle <= 5; cycle++ )
   {
      gmg.solve( L, u, f, maxLevel );
      L.apply( u, r, maxLevel, Inner );
      const real_t res = r.dotGlobal( r, maxLevel, Inner ) / nInner;
      err.assign( { 1.0, -1.0 }, { u, uExact }, maxLevel );
      const real_t e = std::sqrt( err.dotGlobal( err, maxLevel, Inner ) / nInner );
      std::printf( "cycle %d: squared residual %.3e  rate %.3e  error %.3e\n", cycle, res, res / last, e );
      CHECK( res / last < 3.2e-2 || res < 1e-26, "V(3,3) squared-residual reduction worse than the reference's bound 3.2e-2" );
      last = res;
   }

   // Chebyshev smoother as in HyTeG: spectral radius estimate, order-3 polynomial, smoothing reduces the residual
   L.computeInverseDiagonalOperatorValues(); // as in HyTeG: the smoother and the radius estimate need D^-1
   auto cheb = std::make_shared< ChebyshevSmoother< P1ConstantLaplaceOperator > >( storage, minLevel, maxLevel );
   P1Function< real_t > tmp( "tmp", storage, minLevel, maxLevel ), x0( "x0", storage, minLevel, maxLevel );
   x0.interpolate( exact, maxLevel, Inner );
   const real_t rho = chebyshev::estimateRadius( L, maxLevel, 30, storage, x0, tmp );
   std::printf( "spectral radius of D^-1 A: %.4f\n", rho );
   CHECK( rho > 1.0 && rho < 3.0, "spectral radius estimate out of range" );
   cheb->setupCoefficients( 3, rho );
   u.interpolate( real_t( 0 ), maxLevel, Inner );
   L.apply( u, r, maxLevel, Inner );
   const real_t before = r.dotGlobal( r, maxLevel, Inner );
   for ( int i = 0; i < 5; i++ )
      cheb->solve( L, u, f, maxLevel );
   L.apply( u, r, maxLevel, Inner );
   const real_t after = r.dotGlobal( r, maxLevel, Inner );
   std::printf( "Chebyshev(3) x 5: squared residual %.3e -> %.3e\n", before, after );
   CHECK( after < 0.05 * before, "Chebyshev smoothing does not reduce the residual" );

   // error behaviour: src == dst is rejected like in the reference (P1Operator.hpp:198)
   bool threw = false;
   try
   {
      L.apply( u, u, maxLevel, Inner );
   } catch ( const std::exception& )
   {
      threw = true;
   }
   CHECK( threw, "apply( u, u ) must be rejected" );
   std::printf( "OK\n" );
   return 0;
}
