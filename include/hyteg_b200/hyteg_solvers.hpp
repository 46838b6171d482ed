// Operator, smoothers, grid transfer and solver drivers (included by hyteg.hpp).
#pragma once

namespace hyteg {

// ---------------------------------------------------------------------------------------------
// P1ConstantLaplaceOperator  (constant_stencil_operator/P1ConstantOperator.hpp, P1Operator.hpp)
// ---------------------------------------------------------------------------------------------
class P1ConstantLaplaceOperator
{
 public:
   using srcType = P1Function< real_t >;
   using dstType = P1Function< real_t >;

   P1ConstantLaplaceOperator( const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , op_( hbCheckPtr( hb_p1_constant_laplace_create( storage->handle(), (int) minLevel, (int) maxLevel ) ) )
   {}
   /// non-owning view of an existing C-ABI handle
   P1ConstantLaplaceOperator( hb_operator* borrowed, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , op_( borrowed )
   , owned_( false )
   {}
   ~P1ConstantLaplaceOperator()
   {
      if ( owned_ )
         hb_operator_destroy( op_ );
   }
   P1ConstantLaplaceOperator( const P1ConstantLaplaceOperator& )            = delete;
   P1ConstantLaplaceOperator& operator=( const P1ConstantLaplaceOperator& ) = delete;

   std::shared_ptr< PrimitiveStorage > getStorage() const { return storage_; }
   uint_t                              getMinLevel() const { return minLevel_; }
   uint_t                              getMaxLevel() const { return maxLevel_; }
   hb_operator*                        handle() const { return op_; }

   /// Operator::apply (operators/Operator.hpp:76-83)
   void apply( const srcType& src, const dstType& dst, size_t level, DoFType flag, UpdateType updateType = Replace ) const
   {
      hbCheck( hb_operator_apply( op_, src.handle(), dst.handle(), (int) level, (int) flag, (int) updateType ) );
   }
   /// GSSmoothable / SORSmoothable / WeightedJacobiSmoothable (solvers/Smoothables.hpp:34-123)
   void smooth_gs( const srcType& dst, const srcType& rhs, size_t level, DoFType flag ) const { smooth_sor( dst, rhs, 1.0, level, flag ); }
   void smooth_gs_backwards( const srcType& dst, const srcType& rhs, size_t level, DoFType flag ) const { smooth_sor_backwards( dst, rhs, 1.0, level, flag ); }
   void smooth_sor( const srcType& dst, const srcType& rhs, real_t relax, size_t level, DoFType flag ) const
   {
      hbCheck( hb_operator_smooth_sor( op_, dst.handle(), rhs.handle(), relax, (int) level, (int) flag, 0 ) );
   }
   void smooth_sor_backwards( const srcType& dst, const srcType& rhs, real_t relax, size_t level, DoFType flag ) const
   {
      hbCheck( hb_operator_smooth_sor( op_, dst.handle(), rhs.handle(), relax, (int) level, (int) flag, 1 ) );
   }
   /// not in the reference: the sweep with the interior ordering named (0 default, 1 four colours everywhere -- forward and backward
   /// sweeps of ordering 1 are exact transposes of each other; hyteg_b200.h hb_operator_smooth_sor_ordered)
   void smooth_sor_ordered( const srcType& dst, const srcType& rhs, real_t relax, size_t level, DoFType flag, bool backwards, int ordering ) const
   {
      hbCheck( hb_operator_smooth_sor_ordered( op_, dst.handle(), rhs.handle(), relax, (int) level, (int) flag, backwards ? 1 : 0, ordering ) );
   }
   void smooth_jac( const srcType& dst, const srcType& rhs, const srcType& src, real_t relax, size_t level, DoFType flag ) const
   {
      hbCheck( hb_operator_smooth_jac( op_, dst.handle(), rhs.handle(), src.handle(), relax, (int) level, (int) flag ) );
   }
   /// OperatorWithInverseDiagonal
   void computeInverseDiagonalOperatorValues()
   {
      invDiag_ = std::make_shared< srcType >( "inverse diagonal entries", storage_, minLevel_, maxLevel_ );
      for ( uint_t l = minLevel_; l <= maxLevel_; l++ )
         hbCheck( hb_operator_diagonal( op_, invDiag_->handle(), (int) l, 1 ) );
   }
   void computeDiagonalOperatorValues()
   {
      diag_ = std::make_shared< srcType >( "diagonal entries", storage_, minLevel_, maxLevel_ );
      for ( uint_t l = minLevel_; l <= maxLevel_; l++ )
         hbCheck( hb_operator_diagonal( op_, diag_->handle(), (int) l, 0 ) );
   }
   std::shared_ptr< srcType > getInverseDiagonalValues() const
   {
      if ( !invDiag_ )
         throw std::runtime_error( "Inverse diagonal values have not been assembled, call computeInverseDiagonalOperatorValues() to set up this function." );
      return invDiag_;
   }
   std::shared_ptr< srcType > getDiagonalValues() const
   {
      if ( !diag_ )
         throw std::runtime_error( "Diagonal values have not been assembled, call computeDiagonalOperatorValues() to set up this function." );
      return diag_;
   }
   // fused device paths used by the solvers below (no reference counterpart: they replace apply+BLAS-1 chains)
   void residual( const srcType& x, const srcType& b, const srcType& r, size_t level, DoFType flag ) const
   {
      hbCheck( hb_operator_residual( op_, x.handle(), b.handle(), r.handle(), (int) level, (int) flag ) );
   }
   void chebFirst( const srcType& x, const srcType& b, const srcType& dst, size_t level, DoFType flag ) const
   {
      hbCheck( hb_operator_cheb_first( op_, x.handle(), b.handle(), dst.handle(), (int) level, (int) flag ) );
   }
   void chebStep( const srcType& src, const srcType& dst, const srcType& x, real_t coeff, size_t level, DoFType flag ) const
   {
      hbCheck( hb_operator_cheb_step( op_, src.handle(), dst.handle(), x.handle(), coeff, (int) level, (int) flag ) );
   }

 private:
   std::shared_ptr< PrimitiveStorage > storage_;
   uint_t                              minLevel_, maxLevel_;
   hb_operator*                        op_;
   bool                                owned_ = true;
   std::shared_ptr< srcType >          invDiag_, diag_;
};

// ---------------------------------------------------------------------------------------------
// Grid transfer (gridtransferoperators/P1toP1LinearRestriction.hpp / P1toP1LinearProlongation.hpp)
// ---------------------------------------------------------------------------------------------
template < typename FunctionType >
class RestrictionOperator
{
 public:
   virtual ~RestrictionOperator()                                                                       = default;
   virtual void restrict( const FunctionType& function, const uint_t& sourceLevel, const DoFType& flag ) const = 0;
};
template < typename FunctionType >
class ProlongationOperator
{
 public:
   virtual ~ProlongationOperator()                                                                               = default;
   virtual void prolongate( const FunctionType& function, const uint_t& sourceLevel, const DoFType& flag ) const       = 0;
   virtual void prolongateAndAdd( const FunctionType& function, const uint_t& sourceLevel, const DoFType& flag ) const = 0;
};

template < typename ValueType = real_t >
class P1toP1LinearRestriction : public RestrictionOperator< P1Function< ValueType > >
{
 public:
   void restrict( const P1Function< ValueType >& function, const uint_t& sourceLevel, const DoFType& flag ) const override
   {
      hbCheck( hb_p1_restrict( function.handle(), (int) sourceLevel, (int) flag ) );
   }
};

template < typename ValueType = real_t >
class P1toP1LinearProlongation : public ProlongationOperator< P1Function< ValueType > >
{
 public:
   void prolongate( const P1Function< ValueType >& function, const uint_t& sourceLevel, const DoFType& flag ) const override
   {
      hbCheck( hb_p1_prolongate( function.handle(), (int) sourceLevel, (int) flag ) );
   }
   void prolongateAndAdd( const P1Function< ValueType >& function, const uint_t& sourceLevel, const DoFType& flag ) const override
   {
      hbCheck( hb_p1_prolongate_and_add( function.handle(), (int) sourceLevel, (int) flag ) );
   }
};

// ---------------------------------------------------------------------------------------------
// Solvers (solvers/Solver.hpp, CGSolver.hpp, *Smoother.hpp, GeometricMultigridSolver.hpp)
// ---------------------------------------------------------------------------------------------
template < class OperatorType >
class Solver
{
 public:
   using FunctionType = typename OperatorType::srcType;
   virtual ~Solver()  = default;
   virtual void solve( const OperatorType& A, const FunctionType& x, const FunctionType& b, const uint_t level ) = 0;
};

template < class OperatorType >
class IdentityPreconditioner : public Solver< OperatorType >
{
 public:
   void solve( const OperatorType&, const typename OperatorType::srcType& x, const typename OperatorType::srcType& b, const uint_t level ) override
   {
      x.assign( { 1.0 }, { b }, level, All );
   }
};

/// solvers/GaussSeidelSmoother.hpp:30-45
template < class OperatorType >
class GaussSeidelSmoother : public Solver< OperatorType >
{
 public:
   void solve( const OperatorType& A, const typename OperatorType::srcType& x, const typename OperatorType::srcType& b, const uint_t level ) override
   {
      A.smooth_gs( x, b, level, Inner | NeumannBoundary | FreeslipBoundary );
   }
};

/// solvers/SymmetricGaussSeidelSmoother.hpp
template < class OperatorType >
class SymmetricGaussSeidelSmoother : public Solver< OperatorType >
{
 public:
   void solve( const OperatorType& A, const typename OperatorType::srcType& x, const typename OperatorType::srcType& b, const uint_t level ) override
   {
      // the pair must be a forward sweep and its exact transpose (the smoother is used as a symmetric preconditioner as well)
      A.smooth_sor_ordered( x, b, 1.0, level, Inner | NeumannBoundary | FreeslipBoundary, false, 1 );
      A.smooth_sor_ordered( x, b, 1.0, level, Inner | NeumannBoundary | FreeslipBoundary, true, 1 );
   }
};

/// solvers/SORSmoother.hpp:38-50
template < class OperatorType >
class SORSmoother : public Solver< OperatorType >
{
 public:
   explicit SORSmoother( const real_t& relax )
   : relax_( relax )
   {}
   void solve( const OperatorType& A, const typename OperatorType::srcType& x, const typename OperatorType::srcType& b, const uint_t level ) override
   {
      A.smooth_sor( x, b, relax_, level, Inner | NeumannBoundary );
   }

 private:
   real_t relax_;
};

/// solvers/WeightedJacobiSmoother.hpp:36-70
template < class OperatorType >
class WeightedJacobiSmoother : public Solver< OperatorType >
{
 public:
   WeightedJacobiSmoother( const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel, const real_t& relax )
   : relax_( relax )
   , tmp_( "tmp_weighted_jacobi", storage, minLevel, maxLevel )
   , flag_( Inner | NeumannBoundary | FreeslipBoundary )
   {}
   void solve( const OperatorType& A, const typename OperatorType::srcType& x, const typename OperatorType::srcType& b, const uint_t level ) override
   {
      copyBCs( x, tmp_ );
      tmp_.assign( { 1.0 }, { x }, level, All );
      A.smooth_jac( x, b, tmp_, relax_, level, flag_ );
   }

 private:
   real_t                         relax_;
   typename OperatorType::srcType tmp_;
   DoFType                        flag_;
};

/// solvers/ChebyshevSmoother.hpp:112-236 (inverse-diagonal preconditioned Chebyshev polynomial smoother)
template < class OperatorType >
class ChebyshevSmoother : public Solver< OperatorType >
{
 public:
   using FunctionType = typename OperatorType::srcType;

   ChebyshevSmoother( const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , coefficients( maxLevel - minLevel + 1 )
   , tmp1_( "cheb_tmp1", storage, minLevel, maxLevel )
   , tmp2_( "cheb_tmp2", storage, minLevel, maxLevel )
   , flag_( Inner | NeumannBoundary | FreeslipBoundary )
   {}

   void solve( const OperatorType& A, const FunctionType& x, const FunctionType& b, const uint_t level ) override
   {
      const std::vector< real_t >& c = coefficients.at( level - minLevel_ );
      if ( c.empty() )
         throw std::runtime_error( "ChebyshevSmoother: coefficients must be setup" );
      copyBCs( x, tmp1_ );
      copyBCs( x, tmp2_ );
      // tmp1 := D^{-1} (b - A x)                       [reference: apply + assign + multElementwise]
      A.chebFirst( x, b, tmp1_, level, flag_ );
      // k-th step: tmp_new := D^{-1} A tmp_old ; x += c_{k-1} tmp_old   (one fused kernel)
      const FunctionType* cur = &tmp1_;
      const FunctionType* nxt = &tmp2_;
      for ( uint_t k = 1; k < c.size(); k++ )
      {
         A.chebStep( *cur, *nxt, x, c[k - 1], level, flag_ );
         std::swap( cur, nxt );
      }
      x.add( { c.back() }, { *cur }, level, flag_ );
   }

   void setupCoefficients( const uint_t order, const real_t spectralRadius, real_t upperFactor = 1.2, real_t lowerFactor = 0.3 )
   {
      for ( uint_t l = minLevel_; l <= maxLevel_; l++ )
         setupCoefficientsInternal( order, lowerFactor * spectralRadius, upperFactor * spectralRadius, l );
   }
   void setupCoefficients( const uint_t order, const std::vector< real_t >& spectralRadii, real_t upperFactor = 1.2, real_t lowerFactor = 0.3 )
   {
      for ( uint_t l = minLevel_; l <= maxLevel_; l++ )
         setupCoefficientsInternal( order, lowerFactor * spectralRadii.at( l - minLevel_ ), upperFactor * spectralRadii.at( l - minLevel_ ), l );
   }
   void setupCoefficientsOnLevel( const uint_t order, const real_t spectralRadius, const uint_t level, real_t upperFactor = 1.2, real_t lowerFactor = 0.3 )
   {
      setupCoefficientsInternal( order, lowerFactor * spectralRadius, upperFactor * spectralRadius, level );
   }
   /// x_new = x + q(D^-1 A) D^-1 r with residual polynomial 1 - lambda q(lambda) = T_k((theta-lambda)/delta) / T_k(theta/delta),
   /// theta = (ub+lb)/2, delta = (ub-lb)/2: the same polynomial whose monomial coefficients the reference
   /// tabulates in closed form for orders 1..10 (ChebyshevSmoother.hpp:328-539); here expanded by the
   /// three-term recurrence so that any order is available.
   void setupCoefficientsInternal( const uint_t order, const real_t lowerBound, const real_t upperBound, uint_t level )
   {
      if ( order == 0 )
         throw std::runtime_error( "ChebyshevSmoother: order cannot be 0" );
      const long double theta = 0.5L * ( (long double) upperBound + lowerBound );
      const long double delta = 0.5L * ( (long double) upperBound - lowerBound );
      // polynomials in lambda, t(lambda) = a + b*lambda
      const long double                a = theta / delta, bb = -1.0L / delta;
      std::vector< long double >       Tm( 1, 1.0L ), T{ a, bb };
      for ( uint_t n = 1; n < order; n++ )
      {
         std::vector< long double > Tn( T.size() + 1, 0.0L );
         for ( size_t i = 0; i < T.size(); i++ )
         {
            Tn[i] += 2.0L * a * T[i];
            Tn[i + 1] += 2.0L * bb * T[i];
         }
         for ( size_t i = 0; i < Tm.size(); i++ )
            Tn[i] -= Tm[i];
         Tm = T;
         T  = Tn;
      }
      const long double scale = T[0]; // T_k(theta/delta)
      std::vector< real_t >& c = coefficients.at( level - minLevel_ );
      c.assign( order, 0.0 );
      for ( uint_t i = 0; i < order; i++ )
         c[i] = (real_t) ( -T[i + 1] / scale );
   }

   std::vector< std::vector< real_t > > coefficients;

 private:
   uint_t       minLevel_, maxLevel_;
   FunctionType tmp1_, tmp2_;
   DoFType      flag_;
};

/// numerictools/SpectrumEstimation.hpp:56-85 applied to D^-1 A (chebyshev::estimateRadius, ChebyshevSmoother.hpp:655-665)
namespace chebyshev {
template < typename OperatorType >
inline real_t estimateRadiusWithInverseDiagonal( const OperatorType&                   A,
                                                 const typename OperatorType::srcType& invDiag,
                                                 const uint_t&                         level,
                                                 const uint_t&                         maxIter,
                                                 const typename OperatorType::srcType& x,
                                                 const typename OperatorType::srcType& tmp )
{
   auto applyDinvA = [&]( const typename OperatorType::srcType& s, const typename OperatorType::srcType& d ) {
      A.apply( s, d, level, All );
      d.multElementwise( { invDiag, d }, level, All );
   };
   real_t norm = std::sqrt( x.dotGlobal( x, level, All ) );
   x.assign( { 1.0 / norm }, { x }, level, All );
   applyDinvA( x, tmp );
   real_t radius = 0.0;
   for ( uint_t it = 1; it <= maxIter; ++it )
   {
      norm = std::sqrt( tmp.dotGlobal( tmp, level, All ) );
      x.assign( { 1.0 / norm }, { tmp }, level, All );
      applyDinvA( x, tmp );
      radius = x.dotGlobal( tmp, level, All );
   }
   return radius;
}

template < typename OperatorType >
inline real_t estimateRadius( const OperatorType&                        A,
                              const uint_t&                              level,
                              const uint_t&                              maxIter,
                              const std::shared_ptr< PrimitiveStorage >&,
                              typename OperatorType::srcType&            x,
                              typename OperatorType::srcType&            tmp )
{
   return estimateRadiusWithInverseDiagonal( A, *A.getInverseDiagonalValues(), level, maxIter, x, tmp );
}
} // namespace chebyshev

/// solvers/CGSolver.hpp:44-207 (unpreconditioned / preconditioned conjugate gradients)
template < class OperatorType >
class CGSolver : public Solver< OperatorType >
{
 public:
   using FunctionType = typename OperatorType::srcType;

   CGSolver( const std::shared_ptr< PrimitiveStorage >& storage,
             uint_t                                     minLevel,
             uint_t                                     maxLevel,
             uint_t                                     maxIter           = std::numeric_limits< uint_t >::max(),
             real_t                                     relativeTolerance = 1e-16,
             real_t                                     absoluteTolerance = 1e-16,
             std::shared_ptr< Solver< OperatorType > >  preconditioner    = std::make_shared< IdentityPreconditioner< OperatorType > >() )
   : p_( "p", storage, minLevel, maxLevel )
   , z_( "z", storage, minLevel, maxLevel )
   , ap_( "ap", storage, minLevel, maxLevel )
   , r_( "r", storage, minLevel, maxLevel )
   , preconditioner_( preconditioner )
   , flag_( Inner | NeumannBoundary | FreeslipBoundary )
   , printInfo_( false )
   , absoluteTolerance_( absoluteTolerance )
   , relativeTolerance_( relativeTolerance )
   , maxIter_( maxIter )
   , iterations_( maxIter )
   {}

   void setPrintInfo( bool p ) { printInfo_ = p; }
   uint_t getIterations() const { return iterations_; }

   void solve( const OperatorType& A, const FunctionType& x, const FunctionType& b, const uint_t level ) override
   {
      if ( maxIter_ == 0 )
         return;
      copyBCs( x, p_ );
      copyBCs( x, z_ );
      copyBCs( x, ap_ );
      copyBCs( x, r_ );
      p_.setToZero( level );
      z_.setToZero( level );
      ap_.setToZero( level );
      r_.setToZero( level );

      // init (:340-358)
      A.residual( x, b, r_, level, flag_ );
      z_.interpolate( 0.0, level, All );
      preconditioner_->solve( A, z_, r_, level );
      p_.assign( { 1.0 }, { z_ }, level, flag_ );
      real_t prsold = r_.dotGlobal( z_, level, flag_ );

      const real_t res_start = std::sqrt( r_.dotGlobal( r_, level, flag_ ) );
      if ( res_start < absoluteTolerance_ )
      {
         iterations_ = 0;
         return;
      }
      for ( size_t i = 0; i < maxIter_; ++i )
      {
         A.apply( p_, ap_, level, flag_, Replace );
         const real_t pAp   = p_.dotGlobal( ap_, level, flag_ );
         const real_t alpha = prsold / pAp;
         x.add( { alpha }, { p_ }, level, flag_ );
         r_.add( { -alpha }, { ap_ }, level, flag_ );
         const real_t rsnew   = r_.dotGlobal( r_, level, flag_ );
         const real_t sqrsnew = std::sqrt( rsnew );
         const real_t relRes  = sqrsnew / res_start;
         if ( printInfo_ )
            std::cout << "[CG] iter: " << i << ", residual: " << sqrsnew << " ; relative residual: " << relRes << std::endl;
         if ( relRes < relativeTolerance_ || sqrsnew < absoluteTolerance_ )
         {
            iterations_ = i;
            break;
         }
         z_.interpolate( 0.0, level, All );
         preconditioner_->solve( A, z_, r_, level );
         const real_t prsnew = r_.dotGlobal( z_, level, flag_ );
         const real_t beta   = prsnew / prsold;
         p_.assign( { 1.0, beta }, { z_, p_ }, level, flag_ );
         prsold = prsnew;
      }
   }

 private:
   FunctionType                              p_, z_, ap_, r_;
   std::shared_ptr< Solver< OperatorType > > preconditioner_;
   DoFType                                   flag_;
   bool                                      printInfo_;
   real_t                                    absoluteTolerance_, relativeTolerance_;
   uint_t                                    maxIter_, iterations_;
};

/// solvers/GeometricMultigridSolver.hpp:61-312
template < class OperatorType >
class GeometricMultigridSolver : public Solver< OperatorType >
{
 public:
   using FunctionType = typename OperatorType::srcType;

   GeometricMultigridSolver( const std::shared_ptr< PrimitiveStorage >&              storage,
                             std::shared_ptr< Solver< OperatorType > >               smoother,
                             std::shared_ptr< Solver< OperatorType > >               coarseSolver,
                             std::shared_ptr< RestrictionOperator< FunctionType > >  restrictionOperator,
                             std::shared_ptr< ProlongationOperator< FunctionType > > prolongationOperator,
                             uint_t                                                  minLevel,
                             uint_t                                                  maxLevel,
                             uint_t                                                  preSmoothSteps                = 3,
                             uint_t                                                  postSmoothSteps               = 3,
                             uint_t                                                  smoothIncrementOnCoarserGrids = 0,
                             CycleType                                               cycleType = CycleType::VCYCLE )
   : minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , preSmoothSteps_( preSmoothSteps )
   , postSmoothSteps_( postSmoothSteps )
   , smoothIncrement_( smoothIncrementOnCoarserGrids )
   , invokedLevel_( maxLevel )
   , flag_( Inner | NeumannBoundary | FreeslipBoundary )
   , cycleType_( cycleType )
   , smoother_( smoother )
   , coarseSolver_( coarseSolver )
   , restrictionOperator_( restrictionOperator )
   , prolongationOperator_( prolongationOperator )
   , tmp_( std::make_shared< FunctionType >( "gmg_tmp", storage, minLevel, maxLevel ) )
   {}

   void setSmoothingSteps( const uint_t& pre, const uint_t& post, const uint_t smoothIncrement = 0 )
   {
      preSmoothSteps_  = pre;
      postSmoothSteps_ = post;
      smoothIncrement_ = smoothIncrement;
   }

   void solve( const OperatorType& A, const FunctionType& x, const FunctionType& b, const uint_t level ) override
   {
      copyBCs( x, *tmp_ );
      invokedLevel_ = level;
      solveRecursively( A, x, b, level );
   }

   void solveRecursively( const OperatorType& A, const FunctionType& x, const FunctionType& b, const uint_t level ) const
   {
      if ( level == minLevel_ )
      {
         coarseSolver_->solve( A, x, b, minLevel_ );
         return;
      }
      const uint_t pre = preSmoothSteps_ + smoothIncrement_ * ( invokedLevel_ - level );
      for ( uint_t i = 0; i < pre; ++i )
         smoother_->solve( A, x, b, level );

      // tmp = b - A x  (reference: apply + assign, :255-259; fused here)
      A.residual( x, b, *tmp_, level, flag_ );
      restrictionOperator_->restrict( *tmp_, level, flag_ );
      b.assign( { 1.0 }, { *tmp_ }, level - 1, flag_ );
      x.interpolate( 0, level - 1 );

      solveRecursively( A, x, b, level - 1 );
      if ( cycleType_ == CycleType::WCYCLE )
         solveRecursively( A, x, b, level - 1 );

      prolongationOperator_->prolongateAndAdd( x, level - 1, flag_ );

      const uint_t post = postSmoothSteps_ + smoothIncrement_ * ( invokedLevel_ - level );
      for ( uint_t i = 0; i < post; ++i )
         smoother_->solve( A, x, b, level );
   }

 private:
   uint_t    minLevel_, maxLevel_, preSmoothSteps_, postSmoothSteps_, smoothIncrement_, invokedLevel_;
   DoFType   flag_;
   CycleType cycleType_;
   std::shared_ptr< Solver< OperatorType > >               smoother_, coarseSolver_;
   std::shared_ptr< RestrictionOperator< FunctionType > >  restrictionOperator_;
   std::shared_ptr< ProlongationOperator< FunctionType > > prolongationOperator_;
   std::shared_ptr< FunctionType >                         tmp_;
};

} // namespace hyteg
