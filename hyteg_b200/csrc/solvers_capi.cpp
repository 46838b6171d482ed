// C-ABI for the solver drivers: instantiates the C++ host API (include/hyteg_b200/hyteg.hpp), which
// itself only calls the C-ABI of runtime.cu.  Lets Python (ctypes) and other FFI hosts run the same
// GeometricMultigridSolver / CGSolver / smoother code paths as C++ callers.
#include <memory>
#include <string>

#include "../../include/hyteg_b200/hyteg.hpp"
#include "../../include/hyteg_b200/hyteg_p2.hpp"

namespace hb {
void setError( const std::string& s );
}

using Op       = hyteg::P1ConstantLaplaceOperator;
using Fn       = hyteg::P1Function< double >;
using SolverT  = hyteg::Solver< Op >;
using StorageP = std::shared_ptr< hyteg::PrimitiveStorage >;

struct hb_solver
{
   std::shared_ptr< SolverT >                        solver;
   std::shared_ptr< hyteg::ChebyshevSmoother< Op > > cheb;
   std::shared_ptr< hyteg::CGSolver< Op > >          cg;
   StorageP                                          storage;
   int                                               minLevel = 0, maxLevel = 0;
};

#define HB_TRY try {
#define HB_CATCH( ret )              \
   }                                 \
   catch ( const std::exception& e ) \
   {                                 \
      hb::setError( e.what() );      \
      return ret;                    \
   }

static StorageP borrow( hb_storage* st )
{
   return std::make_shared< hyteg::PrimitiveStorage >( st );
}

extern "C" {

hb_solver* hb_solver_create_cg( hb_storage* st, int minLevel, int maxLevel, int maxIter, double relTol, double absTol )
{
   HB_TRY
   auto s     = std::make_unique< hb_solver >();
   s->storage = borrow( st );
   s->cg      = std::make_shared< hyteg::CGSolver< Op > >( s->storage, minLevel, maxLevel, maxIter < 0 ? std::numeric_limits< hyteg::uint_t >::max() : (hyteg::uint_t) maxIter, relTol, absTol );
   s->solver  = s->cg;
   return s.release();
   HB_CATCH( nullptr )
}

hb_solver* hb_solver_create_weighted_jacobi( hb_storage* st, int minLevel, int maxLevel, double omega )
{
   HB_TRY
   auto s     = std::make_unique< hb_solver >();
   s->storage = borrow( st );
   s->solver  = std::make_shared< hyteg::WeightedJacobiSmoother< Op > >( s->storage, minLevel, maxLevel, omega );
   return s.release();
   HB_CATCH( nullptr )
}

hb_solver* hb_solver_create_sor( double omega )
{
   HB_TRY
   auto s    = std::make_unique< hb_solver >();
   s->solver = std::make_shared< hyteg::SORSmoother< Op > >( omega );
   return s.release();
   HB_CATCH( nullptr )
}

hb_solver* hb_solver_create_gauss_seidel( void )
{
   HB_TRY
   auto s    = std::make_unique< hb_solver >();
   s->solver = std::make_shared< hyteg::GaussSeidelSmoother< Op > >();
   return s.release();
   HB_CATCH( nullptr )
}

hb_solver* hb_solver_create_symmetric_gauss_seidel( void )
{
   HB_TRY
   auto s    = std::make_unique< hb_solver >();
   s->solver = std::make_shared< hyteg::SymmetricGaussSeidelSmoother< Op > >();
   return s.release();
   HB_CATCH( nullptr )
}

hb_solver* hb_solver_create_chebyshev( hb_storage* st, int minLevel, int maxLevel )
{
   HB_TRY
   auto s      = std::make_unique< hb_solver >();
   s->storage  = borrow( st );
   s->cheb     = std::make_shared< hyteg::ChebyshevSmoother< Op > >( s->storage, minLevel, maxLevel );
   s->solver   = s->cheb;
   s->minLevel = minLevel;
   s->maxLevel = maxLevel;
   return s.release();
   HB_CATCH( nullptr )
}

int hb_chebyshev_setup_coefficients( hb_solver* s, int order, double lb, double ub )
{
   HB_TRY
   if ( !s->cheb )
      throw std::runtime_error( "hyteg_b200: not a Chebyshev smoother" );
   for ( int l = s->minLevel; l <= s->maxLevel; l++ )
      s->cheb->setupCoefficientsInternal( order, lb, ub, l );
   return 0;
   HB_CATCH( 1 )
}

int hb_chebyshev_coefficients( const hb_solver* s, double* coeffs, int* n )
{
   HB_TRY
   if ( !s->cheb )
      throw std::runtime_error( "hyteg_b200: not a Chebyshev smoother" );
   const auto& c = s->cheb->coefficients.at( 0 );
   *n            = (int) c.size();
   for ( size_t i = 0; i < c.size(); i++ )
      coeffs[i] = c[i];
   return 0;
   HB_CATCH( 1 )
}

int hb_estimate_spectral_radius( hb_operator* op, hb_function* x, hb_function* tmp, int iterations, int level, double* radius )
{
   HB_TRY
   int a, b, fa, fb;
   hb_operator_level_range( op, &a, &b );
   hb_p1function_level_range( x, &fa, &fb );
   StorageP st = borrow( hb_p1function_storage( x ) );
   Op       A( op, st, a, b );
   Fn       X( x, st, fa, fb );
   hb_p1function_level_range( tmp, &fa, &fb );
   Fn T( tmp, st, fa, fb );
   Fn invDiag( "inverse diagonal entries", st, level, level );
   hyteg::hbCheck( hb_operator_diagonal( op, invDiag.handle(), level, 1 ) );
   *radius = hyteg::chebyshev::estimateRadiusWithInverseDiagonal( A, invDiag, level, iterations, X, T );
   return 0;
   HB_CATCH( 1 )
}

hb_solver* hb_solver_create_gmg( hb_storage* st, hb_solver* smoother, hb_solver* coarse, int minLevel, int maxLevel, int pre, int post, int inc, int cycleType )
{
   HB_TRY
   auto s     = std::make_unique< hb_solver >();
   s->storage = borrow( st );
   s->solver  = std::make_shared< hyteg::GeometricMultigridSolver< Op > >( s->storage,
                                                                         smoother->solver,
                                                                         coarse->solver,
                                                                         std::make_shared< hyteg::P1toP1LinearRestriction<> >(),
                                                                         std::make_shared< hyteg::P1toP1LinearProlongation<> >(),
                                                                         minLevel,
                                                                         maxLevel,
                                                                         pre,
                                                                         post,
                                                                         inc,
                                                                         cycleType == HB_WCYCLE ? hyteg::CycleType::WCYCLE : hyteg::CycleType::VCYCLE );
   return s.release();
   HB_CATCH( nullptr )
}

int hb_solver_solve( hb_solver* s, hb_operator* op, hb_function* x, hb_function* b, int level )
{
   HB_TRY
   int a, bb, fa, fb;
   hb_operator_level_range( op, &a, &bb );
   hb_p1function_level_range( x, &fa, &fb );
   Op A( op, s->storage, a, bb );
   Fn X( x, s->storage, fa, fb );
   hb_p1function_level_range( b, &fa, &fb );
   Fn B( b, s->storage, fa, fb );
   s->solver->solve( A, X, B, (hyteg::uint_t) level );
   return 0;
   HB_CATCH( 1 )
}

/* ---- P2: Chebyshev-smoothed GMG on an elementwise operator (the recipe of src/terraneo/solvers/StokesMCFGMRESSolver.hpp:71-143 for the
 * velocity block: GeometricMultigridSolver + ChebyshevSmoother( inverse diagonal ) + quadratic transfer; CG on the coarsest level) ---- */
struct hb_p2solver
{
   using P2Op = hyteg::P2ElementwiseOperatorB200;
   StorageP                                                       storage;
   std::shared_ptr< P2Op >                                        A;
   std::shared_ptr< hyteg::ChebyshevSmoother< P2Op > >            cheb;
   std::shared_ptr< hyteg::GeometricMultigridSolver< P2Op > >     gmg;
   double                                                         rho = 0;
   int                                                            minLevel = 0, maxLevel = 0;
};

hb_p2solver* hb_p2_gmg_chebyshev_create( hb_storage* st, hb_p2operator* op, int minLevel, int maxLevel, int order, hb_p2function* eigenvector_guess,
                                         int powerIterations, int pre, int post, int coarseMaxIter, double coarseTol )
{
   HB_TRY
   using P2Op = hb_p2solver::P2Op;
   using P2Fn = hyteg::P2Function< double >;
   auto s      = std::make_unique< hb_p2solver >();
   s->storage  = borrow( st );
   s->minLevel = minLevel;
   s->maxLevel = maxLevel;
   s->A        = std::make_shared< P2Op >( op, s->storage, (hyteg::uint_t) minLevel, (hyteg::uint_t) maxLevel );
   s->A->computeInverseDiagonalOperatorValues();
   P2Fn ev( eigenvector_guess, s->storage, (hyteg::uint_t) maxLevel, (hyteg::uint_t) maxLevel );
   P2Fn tmp( "tmp", s->storage, (hyteg::uint_t) maxLevel, (hyteg::uint_t) maxLevel );
   s->rho  = hyteg::chebyshev::estimateRadius( *s->A, (hyteg::uint_t) maxLevel, (hyteg::uint_t) powerIterations, s->storage, ev, tmp );
   s->cheb = std::make_shared< hyteg::ChebyshevSmoother< P2Op > >( s->storage, minLevel, maxLevel );
   s->cheb->setupCoefficients( (hyteg::uint_t) order, s->rho );
   auto coarse = std::make_shared< hyteg::CGSolver< P2Op > >( s->storage, minLevel, minLevel, (hyteg::uint_t) coarseMaxIter, coarseTol, coarseTol );
   s->gmg      = std::make_shared< hyteg::GeometricMultigridSolver< P2Op > >( s->storage, s->cheb, coarse, std::make_shared< hyteg::P2toP2QuadraticRestriction >(),
                                                                              std::make_shared< hyteg::P2toP2QuadraticProlongation >(), minLevel, maxLevel, pre, post );
   return s.release();
   HB_CATCH( nullptr )
}

double hb_p2solver_spectral_radius( const hb_p2solver* s )
{
   return s->rho;
}

/* what = 0: one V-cycle (GeometricMultigridSolver::solve), 1: one call of the Chebyshev smoother on 'level' */
int hb_p2solver_solve( hb_p2solver* s, hb_p2function* x, hb_p2function* b, int level, int what )
{
   HB_TRY
   using P2Fn = hyteg::P2Function< double >;
   P2Fn X( x, s->storage, (hyteg::uint_t) s->minLevel, (hyteg::uint_t) s->maxLevel ), B( b, s->storage, (hyteg::uint_t) s->minLevel, (hyteg::uint_t) s->maxLevel );
   if ( what == 1 )
      s->cheb->solve( *s->A, X, B, (hyteg::uint_t) level );
   else
      s->gmg->solve( *s->A, X, B, (hyteg::uint_t) level );
   return 0;
   HB_CATCH( 1 )
}

void hb_p2solver_destroy( hb_p2solver* s )
{
   delete s;
}

int hb_solver_iterations( const hb_solver* s )
{
   return s->cg ? (int) s->cg->getIterations() : -1;
}

void hb_solver_destroy( hb_solver* s )
{
   delete s;
}

} // extern "C"
