// C-ABI for the solver drivers: instantiates the C++ host API (include/hyteg_b200/hyteg.hpp), which
// itself only calls the C-ABI of runtime.cu.  Lets Python (ctypes) and other FFI hosts run the same
// GeometricMultigridSolver / CGSolver / smoother code paths as C++ callers.
#include <cmath>
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

// Direct solver for level 0 of an UNPARTITIONED storage (the reference's counterpart as a coarse-grid solver is PETScLUSolver,
// src/hyteg/petsc/PETScLUSolver.hpp): the unknowns of level 0 are the macro vertices, a few hundred at most on the meshes this library is
// run on.  The dense matrix of the free DoFs is assembled once per boundary-condition table by applying the operator to unit vectors,
// Cholesky-factorised on the host, and a solve is one small download, two triangular sweeps and one upload -- instead of 40-50 CG
// iterations of ~10 launches and three host-synchronising dot products each.
class DirectLevel0Solver : public SolverT
{
 public:
   explicit DirectLevel0Solver( hb_storage* st )
   : st_( st )
   {
      int counts[4];
      hyteg::hbCheck( hb_storage_counts( st, counts ) );
      nV_ = counts[0];
      nC_ = hb_storage_num_local_cells( st );
      if ( nC_ != counts[3] )
         throw std::runtime_error( "hyteg_b200: the direct level-0 solver needs an unpartitioned storage (use it on a replica)" );
      cellV_.resize( (size_t) 4 * nC_ );
      hyteg::hbCheck( hb_storage_primitive_vertices( st, 3, cellV_.data() ) );
   }
   void solve( const Op& A, const Fn& x, const Fn& b, const hyteg::uint_t level ) override
   {
      if ( level != 0 )
         throw std::runtime_error( "hyteg_b200: the direct coarse solver works on level 0" );
      const int flag = hyteg::Inner | hyteg::NeumannBoundary | hyteg::FreeslipBoundary;
      if ( !tmp_ )
      {
         tmp_  = std::make_shared< Fn >( "direct tmp", x.getStorage(), 0, 0 );
         tmp2_ = std::make_shared< Fn >( "direct tmp2", x.getStorage(), 0, 0 );
      }
      std::vector< int > bc( 16 );
      hyteg::hbCheck( hb_p1function_get_boundary_condition( x.handle(), bc.data(), 16 ) );
      if ( bc != bc_ )
         factorise( A, x, bc, flag );
      std::vector< double > hx = gatherVertices( x ), hb = gatherVertices( b );
      // rhs of the free rows: b_f - A_fd x_d
      const int             n = (int) free_.size();
      std::vector< double > y( (size_t) n );
      for ( int i = 0; i < n; i++ )
      {
         double s = hb[(size_t) free_[(size_t) i]];
         for ( int j = 0; j < nV_; j++ )
            if ( !isFree_[(size_t) j] )
               s -= rows_[(size_t) i * nV_ + j] * hx[(size_t) j];
         y[(size_t) i] = s;
      }
      for ( int i = 0; i < n; i++ ) // L z = y
      {
         double s = y[(size_t) i];
         for ( int k = 0; k < i; k++ )
            s -= chol_[(size_t) i * n + k] * y[(size_t) k];
         y[(size_t) i] = s / chol_[(size_t) i * n + i];
      }
      for ( int i = n - 1; i >= 0; i-- ) // L^T x = z
      {
         double s = y[(size_t) i];
         for ( int k = i + 1; k < n; k++ )
            s -= chol_[(size_t) k * n + i] * y[(size_t) k];
         y[(size_t) i] = s / chol_[(size_t) i * n + i];
      }
      for ( int i = 0; i < n; i++ )
         hx[(size_t) free_[(size_t) i]] = y[(size_t) i];
      scatterVertices( x, hx, flag );
   }

 private:
   std::vector< double > gatherVertices( const Fn& f ) const
   {
      std::vector< double > cells( (size_t) 4 * nC_ ), v( (size_t) nV_, 0.0 );
      hyteg::hbCheck( hb_p1function_download( f.handle(), 0, cells.data() ) );
      for ( size_t i = 0; i < cells.size(); i++ )
         v[(size_t) cellV_[i]] = cells[i]; // all copies of a macro vertex hold the same value
      return v;
   }
   void scatterVertices( const Fn& f, const std::vector< double >& v, int flag ) const
   {
      std::vector< double > cells( (size_t) 4 * nC_ );
      for ( size_t i = 0; i < cells.size(); i++ )
         cells[i] = v[(size_t) cellV_[i]];
      hyteg::hbCheck( hb_p1function_upload( f.handle(), 0, cells.data(), flag ) );
   }
   void factorise( const Op& A, const Fn& x, const std::vector< int >& bc, int flag )
   {
      hyteg::copyBCs( x, *tmp_ );
      hyteg::copyBCs( x, *tmp2_ );
      tmp_->interpolate( 0.0, 0, hyteg::All );
      tmp_->interpolate( 1.0, 0, (hyteg::DoFType) flag );
      const std::vector< double > mask = gatherVertices( *tmp_ );
      isFree_.assign( (size_t) nV_, 0 );
      free_.clear();
      for ( int j = 0; j < nV_; j++ )
         if ( mask[(size_t) j] != 0.0 )
         {
            isFree_[(size_t) j] = 1;
            free_.push_back( j );
         }
      const int n = (int) free_.size();
      rows_.assign( (size_t) n * nV_, 0.0 );
      std::vector< double > e( (size_t) nV_, 0.0 );
      for ( int j = 0; j < nV_; j++ ) // column j of the free rows: A e_j
      {
         e[(size_t) j] = 1.0;
         scatterVertices( *tmp_, e, hyteg::All );
         e[(size_t) j] = 0.0;
         tmp2_->interpolate( 0.0, 0, hyteg::All );
         A.apply( *tmp_, *tmp2_, 0, (hyteg::DoFType) flag, hyteg::Replace );
         const std::vector< double > col = gatherVertices( *tmp2_ );
         for ( int i = 0; i < n; i++ )
            rows_[(size_t) i * nV_ + j] = col[(size_t) free_[(size_t) i]];
      }
      chol_.assign( (size_t) n * n, 0.0 ); // A_ff = L L^T, lower triangle, row-major
      for ( int i = 0; i < n; i++ )
         for ( int k = 0; k <= i; k++ )
         {
            double s = rows_[(size_t) i * nV_ + free_[(size_t) k]];
            for ( int m = 0; m < k; m++ )
               s -= chol_[(size_t) i * n + m] * chol_[(size_t) k * n + m];
            if ( i == k )
            {
               if ( !( s > 0.0 ) )
                  throw std::runtime_error( "hyteg_b200: the level-0 operator is not positive definite on the free DoFs (direct coarse solver)" );
               chol_[(size_t) i * n + i] = std::sqrt( s );
            }
            else
               chol_[(size_t) i * n + k] = s / chol_[(size_t) k * n + k];
         }
      bc_ = bc;
   }

   hb_storage*           st_;
   int                   nV_ = 0, nC_ = 0;
   std::vector< int >    cellV_, free_, bc_;
   std::vector< char >   isFree_;
   std::vector< double > rows_, chol_;
   std::shared_ptr< Fn > tmp_, tmp2_;
};

// Coarse part of a multigrid cycle on a REPLICA of a partitioned storage: as the coarse-grid solver of a GeometricMultigridSolver on
// levels coarseLevel .. L it runs, on every rank redundantly, exactly what the cycle would do on levels 0 .. coarseLevel -- pre-smoothing,
// restriction, ..., CG on level 0, ..., post-smoothing -- after one all-gather of x and b and before a local copy back.  The levels
// below coarseLevel then cost no ghost exchange at all (a partitioned V(3,3) cycle pays about 40 of them per level, ~10 us each, and
// three all-reduces per CG iteration).
class ReplicatedCoarseCycle : public SolverT
{
 public:
   ReplicatedCoarseCycle( hb_storage* part, int coarseLevel, int smoother, double omega, int pre, int post, int cgMaxIter, double cgRel, double cgAbs )
   : level_( (hyteg::uint_t) coarseLevel )
   {
      hb_storage* r = hb_storage_create_replica( part );
      if ( !r )
         throw std::runtime_error( hb_last_error() );
      rep_ = std::shared_ptr< hyteg::PrimitiveStorage >( new hyteg::PrimitiveStorage( r ), [r]( hyteg::PrimitiveStorage* p ) {
         delete p;
         hb_storage_destroy( r );
      } );
      A_ = std::make_shared< Op >( rep_, 0, level_ );
      x_ = std::make_shared< Fn >( "replica x", rep_, 0, level_ );
      b_ = std::make_shared< Fn >( "replica b", rep_, 0, level_ );
      std::shared_ptr< SolverT > sm;
      if ( smoother == 1 )
         sm = std::make_shared< hyteg::WeightedJacobiSmoother< Op > >( rep_, 0, level_, omega );
      else
         sm = std::make_shared< hyteg::GaussSeidelSmoother< Op > >();
      std::shared_ptr< SolverT > cg; // cgMaxIter == 0: direct solve on level 0
      if ( cgMaxIter == 0 )
         cg = std::make_shared< DirectLevel0Solver >( r );
      else
         cg = std::make_shared< hyteg::CGSolver< Op > >( rep_, 0, 0, cgMaxIter < 0 ? std::numeric_limits< hyteg::uint_t >::max() : (hyteg::uint_t) cgMaxIter, cgRel, cgAbs );
      if ( coarseLevel == 0 )
         cycle_ = cg;
      else
         cycle_ = std::make_shared< hyteg::GeometricMultigridSolver< Op > >( rep_,
                                                                             sm,
                                                                             cg,
                                                                             std::make_shared< hyteg::P1toP1LinearRestriction<> >(),
                                                                             std::make_shared< hyteg::P1toP1LinearProlongation<> >(),
                                                                             0,
                                                                             level_,
                                                                             pre,
                                                                             post );
   }
   void solve( const Op&, const Fn& x, const Fn& b, const hyteg::uint_t level ) override
   {
      if ( level != level_ )
         throw std::runtime_error( "hyteg_b200: the replicated coarse cycle was built for level " + std::to_string( level_ ) + ", called on level " + std::to_string( level ) );
      hyteg::copyBCs( x, *x_ );
      hyteg::copyBCs( b, *b_ );
      hyteg::hbCheck( hb_p1function_gather_to_replica( x.handle(), x_->handle(), (int) level ) );
      hyteg::hbCheck( hb_p1function_gather_to_replica( b.handle(), b_->handle(), (int) level ) );
      cycle_->solve( *A_, *x_, *b_, level );
      hyteg::hbCheck( hb_p1function_scatter_from_replica( x.handle(), x_->handle(), (int) level ) );
   }

 private:
   hyteg::uint_t               level_;
   StorageP                    rep_;
   std::shared_ptr< Op >       A_;
   std::shared_ptr< Fn >       x_, b_;
   std::shared_ptr< SolverT >  cycle_;
};

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

hb_solver* hb_solver_create_replicated_coarse_cycle( hb_storage* st, int coarseLevel, int smoother, double omega, int pre, int post, int cgMaxIter, double cgRelTol,
                                                     double cgAbsTol )
{
   HB_TRY
   auto s     = std::make_unique< hb_solver >();
   s->storage = borrow( st );
   s->solver  = std::make_shared< ReplicatedCoarseCycle >( st, coarseLevel, smoother, omega, pre, post, cgMaxIter, cgRelTol, cgAbsTol );
   return s.release();
   HB_CATCH( nullptr )
}

hb_solver* hb_solver_create_direct_level0( hb_storage* st )
{
   HB_TRY
   auto s     = std::make_unique< hb_solver >();
   s->storage = borrow( st );
   s->solver  = std::make_shared< DirectLevel0Solver >( st );
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
