// hyteg_b200 -- C++ mirror of HyTeG's P2 interfaces for the fused elementwise path, header-only over the C-ABI (hyteg_b200.h).
//
//   geometry/IcosahedralShellMap.hpp:491-510                IcosahedralShellMap::setMap
//   p2functionspace/P2Function.hpp:40-300                   P2Function (vertexdof + edgedof, :76-77)
//   elementwiseoperators/P2ElementwiseOperator.hpp:63-100   P2ElementwiseOperator< Form >::apply / gemv
//   elementwiseoperators/P2ElementwiseOperator.hpp:475      P2ElementwiseBlendingDivKGradOperator
//   elementwiseoperators/P2ElementwiseEpsilonOperator.hpp:112-170  the nine blocks of the epsilon operator
// Same names, argument meaning and error behaviour (exceptions) as the reference; all arithmetic runs in the CUDA library.
#pragma once
#include "hyteg.hpp"

namespace hyteg {

struct IcosahedralShellMap
{
   /// IcosahedralShellMap::setMap( setupStorage ): here on the built storage (the map parameters live with the device tables)
   static void setMap( PrimitiveStorage& storage ) { hbCheck( hb_storage_set_icosahedral_shell_map( storage.handle() ) ); }
};

template < typename ValueType = real_t >
class P2Function
{
   static_assert( std::is_same< ValueType, double >::value, "the B200 path is FP64" );

 public:
   using valueType = ValueType;

   P2Function( const std::string& name, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : name_( name )
   , storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , f_( hbCheckPtr( hb_p2function_create( storage->handle(), (int) minLevel, (int) maxLevel ) ) )
   , v_( hb_p2function_vertex_part( f_ ), storage, minLevel, maxLevel )
   {}
   /// non-owning view of an existing C-ABI handle
   P2Function( hb_p2function* borrowed, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : name_( "borrowed" )
   , storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , f_( borrowed )
   , v_( hb_p2function_vertex_part( f_ ), storage, minLevel, maxLevel )
   , owned_( false )
   {}
   ~P2Function()
   {
      if ( owned_ )
         hb_p2function_destroy( f_ );
   }
   P2Function( const P2Function& )            = delete;
   P2Function& operator=( const P2Function& ) = delete;

   const std::string&                  getFunctionName() const { return name_; }
   std::shared_ptr< PrimitiveStorage > getStorage() const { return storage_; }
   uint_t                              getMinLevel() const { return minLevel_; }
   uint_t                              getMaxLevel() const { return maxLevel_; }
   hb_p2function*                      handle() const { return f_; }
   /// P2Function::getVertexDoFFunction (P2Function.hpp:76)
   const P1Function< ValueType >& getVertexDoFFunction() const { return v_; }

   void interpolate( ValueType constant, uint_t level, DoFType flag = All ) const
   {
      hbCheck( hb_p2function_interpolate_constant( f_, constant, (int) level, (int) flag ) );
   }
   /// expr is evaluated at the (blended, if a geometry map is set) position of every vertex DoF and every edge midpoint; all DoFs are
   /// written (flag All), as a test set-up would do
   void interpolate( const std::function< ValueType( const Point3D& ) >& expr, uint_t level, bool blended = false ) const
   {
      hb_storage*  st = storage_->handle();
      const int    nc = hb_storage_num_local_cells( st );
      const size_t nv = (size_t) hb_cell_array_size( (int) level ), ne = (size_t) hb_p2_edge_cell_size( (int) level );
      std::vector< double > xyz( 3 * nv * (size_t) nc ), val( nv * (size_t) nc ), ev( ne * (size_t) nc ), ends( 6 * ne ), mid( 3 * ne ), F( 12 * std::max( ne, nv ) );
      hbCheck( hb_storage_micro_coordinates( st, (int) level, xyz.data() ) );
      for ( int c = 0; c < nc; c++ )
      {
         double* x = xyz.data() + 3 * nv * (size_t) c;
         if ( blended )
         {
            hbCheck( hb_storage_map_eval( st, (int) level, c, x, (int) nv, F.data() ) );
            for ( size_t i = 0; i < nv; i++ )
               for ( int d = 0; d < 3; d++ )
                  x[3 * i + (size_t) d] = F[12 * i + (size_t) d];
         }
         for ( size_t i = 0; i < nv; i++ )
            val[nv * (size_t) c + i] = expr( Point3D( x[3 * i], x[3 * i + 1], x[3 * i + 2] ) );
         hbCheck( hb_p2_edge_endpoints( st, (int) level, c, ends.data() ) );
         for ( size_t i = 0; i < ne; i++ )
            for ( int d = 0; d < 3; d++ )
               mid[3 * i + (size_t) d] = 0.5 * ( ends[6 * i + (size_t) d] + ends[6 * i + 3 + (size_t) d] );
         if ( blended )
         {
            hbCheck( hb_storage_map_eval( st, (int) level, c, mid.data(), (int) ne, F.data() ) );
            for ( size_t i = 0; i < ne; i++ )
               for ( int d = 0; d < 3; d++ )
                  mid[3 * i + (size_t) d] = F[12 * i + (size_t) d];
         }
         for ( size_t i = 0; i < ne; i++ )
            ev[ne * (size_t) c + i] = expr( Point3D( mid[3 * i], mid[3 * i + 1], mid[3 * i + 2] ) );
      }
      hbCheck( hb_p1function_upload( v_.handle(), (int) level, val.data(), (int) All ) );
      hbCheck( hb_p2function_upload_edges( f_, (int) level, ev.data() ) );
   }
   void assign( const std::vector< ValueType >& scalars, const std::vector< std::reference_wrapper< const P2Function > >& functions, uint_t level, DoFType flag = All ) const
   {
      std::vector< hb_p2function* > h;
      for ( const P2Function& f : functions )
         h.push_back( f.f_ );
      hbCheck( hb_p2function_assign( f_, (int) h.size(), scalars.data(), h.data(), (int) level, (int) flag ) );
   }
   void add( const std::vector< ValueType >& scalars, const std::vector< std::reference_wrapper< const P2Function > >& functions, uint_t level, DoFType flag = All ) const
   {
      std::vector< hb_p2function* > h;
      for ( const P2Function& f : functions )
         h.push_back( f.f_ );
      hbCheck( hb_p2function_add( f_, (int) h.size(), scalars.data(), h.data(), (int) level, (int) flag ) );
   }
   ValueType dotGlobal( const P2Function& rhs, uint_t level, DoFType flag = All ) const
   {
      double v;
      hbCheck( hb_p2function_dot( f_, rhs.f_, (int) level, (int) flag, &v ) );
      return v;
   }
   void multElementwise( const std::vector< std::reference_wrapper< const P2Function > >& functions, uint_t level, DoFType flag = All ) const
   {
      std::vector< hb_p2function* > h;
      for ( const P2Function& f : functions )
         h.push_back( f.f_ );
      hbCheck( hb_p2function_mult_elementwise( f_, (int) h.size(), h.data(), (int) level, (int) flag ) );
   }
   void setToZero( uint_t level ) const { interpolate( ValueType( 0 ), level, All ); }
   void copyBoundaryConditionFromFunction( const P2Function& other ) const { hbCheck( hb_p2function_copy_boundary_condition( f_, other.f_ ) ); }

 private:
   std::string                         name_;
   std::shared_ptr< PrimitiveStorage > storage_;
   uint_t                              minLevel_, maxLevel_;
   hb_p2function*                      f_;
   P1Function< ValueType >             v_; // non-owning view of the vertex part
   bool                                owned_ = true;
};

/// P2ElementwiseOperator< Form > for the forms of the fused path; the coefficient is a P2Function (the hyteg_operators interface), a host
/// callback sampled at the blended quadrature points (the in-tree forms' std::function), or absent (= 1)
class P2ElementwiseOperatorB200
{
 public:
   using srcType = P2Function< real_t >;
   using dstType = P2Function< real_t >;
   /// non-owning view of an existing C-ABI handle
   P2ElementwiseOperatorB200( hb_p2operator* borrowed, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : op_( borrowed )
   , owned_( false )
   , storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   {}
   virtual ~P2ElementwiseOperatorB200()
   {
      if ( owned_ )
         hb_p2operator_destroy( op_ );
   }
   void apply( const srcType& src, const dstType& dst, size_t level, DoFType flag, UpdateType updateType = Replace ) const
   {
      hbCheck( hb_p2operator_apply( op_, src.handle(), dst.handle(), (int) level, (int) flag, (int) updateType ) );
   }
   hb_p2operator* handle() const { return op_; }
   /// OperatorWithInverseDiagonal (P2ElementwiseOperator.hpp:133-160)
   void computeInverseDiagonalOperatorValues()
   {
      invDiag_ = std::make_shared< srcType >( "inverse diagonal entries", storage_, minLevel_, maxLevel_ );
      for ( uint_t l = minLevel_; l <= maxLevel_; l++ )
      {
         hb_p2function* f = invDiag_->handle();
         hbCheck( hb_p2operator_diagonal( op_, &f, (int) l, 1 ) );
      }
   }
   std::shared_ptr< srcType > getInverseDiagonalValues() const
   {
      if ( !invDiag_ )
         throw std::runtime_error( "Inverse diagonal values have not been assembled, call computeInverseDiagonalOperatorValues() to set up this function." );
      return invDiag_;
   }
   // the building blocks the solver templates (hyteg_solvers.hpp) call; here chains of apply + BLAS-1 like in the reference
   void residual( const srcType& x, const srcType& b, const srcType& r, size_t level, DoFType flag ) const
   {
      apply( x, r, level, flag, Replace );
      r.assign( { 1.0, -1.0 }, { b, r }, level, flag );
   }
   void chebFirst( const srcType& x, const srcType& b, const srcType& dst, size_t level, DoFType flag ) const
   {
      residual( x, b, dst, level, flag );
      dst.multElementwise( { *getInverseDiagonalValues(), dst }, level, flag );
   }
   void chebStep( const srcType& src, const srcType& dst, const srcType& x, real_t coeff, size_t level, DoFType flag ) const
   {
      apply( src, dst, level, flag, Replace );
      dst.multElementwise( { *getInverseDiagonalValues(), dst }, level, flag );
      x.add( { coeff }, { src }, level, flag );
   }

 protected:
   P2ElementwiseOperatorB200( hb_p2operator* op, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel,
                              const std::function< real_t( const Point3D& ) >* k )
   : op_( hbCheckPtr( op ) )
   , storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   {
      if ( k )
         for ( uint_t L = minLevel; L <= maxLevel; L++ )
         {
            const size_t          n = (size_t) hb_p2operator_num_quadrature_points( op_, (int) L );
            std::vector< double > xyz( 3 * n ), val( n );
            hbCheck( hb_p2operator_quadrature_points( op_, (int) L, xyz.data() ) );
            for ( size_t i = 0; i < n; i++ )
               val[i] = ( *k )( Point3D( xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2] ) );
            hbCheck( hb_p2operator_set_sampled_coefficient( op_, (int) L, val.data() ) );
         }
   }
   hb_p2operator*                      op_;
   bool                                owned_ = true;
   std::shared_ptr< PrimitiveStorage > storage_;
   uint_t                              minLevel_ = 0, maxLevel_ = 0;
   std::shared_ptr< srcType >          invDiag_;
};

/// P2ElementwiseBlendingDivKGradOperator (P2ElementwiseOperator.hpp:475) / operatorgeneration::P2ElementwiseDivKGrad...
class P2ElementwiseBlendingDivKGradOperator : public P2ElementwiseOperatorB200
{
 public:
   P2ElementwiseBlendingDivKGradOperator( const std::shared_ptr< PrimitiveStorage >& storage, size_t minLevel, size_t maxLevel, const P2Function< real_t >& k )
   : P2ElementwiseOperatorB200( hb_p2_divkgrad_create( storage->handle(), (int) minLevel, (int) maxLevel, k.handle() ), storage, minLevel, maxLevel, nullptr )
   {}
   P2ElementwiseBlendingDivKGradOperator( const std::shared_ptr< PrimitiveStorage >& storage, size_t minLevel, size_t maxLevel, const std::function< real_t( const Point3D& ) >& k )
   : P2ElementwiseOperatorB200( hb_p2_divkgrad_create( storage->handle(), (int) minLevel, (int) maxLevel, nullptr ), storage, minLevel, maxLevel, &k )
   {}
};

/// block (i, j) of P2ElementwiseBlendingEpsilonOperator (P2ElementwiseEpsilonOperator.hpp:112-170)
class P2ElementwiseBlendingEpsilonBlock : public P2ElementwiseOperatorB200
{
 public:
   P2ElementwiseBlendingEpsilonBlock( const std::shared_ptr< PrimitiveStorage >& storage, size_t minLevel, size_t maxLevel, int i, int j, const std::function< real_t( const Point3D& ) >& mu )
   : P2ElementwiseOperatorB200( hb_p2_epsilon_create( storage->handle(), (int) minLevel, (int) maxLevel, i, j, nullptr ), storage, minLevel, maxLevel, &mu )
   {}
   P2ElementwiseBlendingEpsilonBlock( const std::shared_ptr< PrimitiveStorage >& storage, size_t minLevel, size_t maxLevel, int i, int j, const P2Function< real_t >& mu )
   : P2ElementwiseOperatorB200( hb_p2_epsilon_create( storage->handle(), (int) minLevel, (int) maxLevel, i, j, mu.handle() ), storage, minLevel, maxLevel, nullptr )
   {}
};

/// gridtransferoperators/P2toP2QuadraticRestriction.hpp / P2toP2QuadraticProlongation.hpp
class P2toP2QuadraticRestriction : public RestrictionOperator< P2Function< real_t > >
{
 public:
   void restrict( const P2Function< real_t >& function, const uint_t& sourceLevel, const DoFType& flag ) const override
   {
      hbCheck( hb_p2_restrict( function.handle(), (int) sourceLevel, (int) flag ) );
   }
};
class P2toP2QuadraticProlongation : public ProlongationOperator< P2Function< real_t > >
{
 public:
   void prolongate( const P2Function< real_t >& function, const uint_t& sourceLevel, const DoFType& flag ) const override
   {
      hbCheck( hb_p2_prolongate( function.handle(), (int) sourceLevel, (int) flag ) );
   }
   void prolongateAndAdd( const P2Function< real_t >& function, const uint_t& sourceLevel, const DoFType& flag ) const override
   {
      hbCheck( hb_p2_prolongate_and_add( function.handle(), (int) sourceLevel, (int) flag ) );
   }
};

} // namespace hyteg
