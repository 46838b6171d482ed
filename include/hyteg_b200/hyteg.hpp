// C++ host API of hyteg_b200: HyTeG's own class names and call signatures for the matrix-free P1
// multigrid path, implemented as thin wrappers over the C-ABI (include/hyteg_b200.h).  Host code that
// is written against HyTeG's Operator / Function / Solver interfaces for this path compiles against
// this header by swapping the include; all arithmetic runs in the sm_100a kernels behind the C-ABI.
//
// Mirrored reference interfaces (paths relative to the HyTeG tree):
//   types/types.hpp:29-97                         DoFType, UpdateType, CycleType, testFlag
//   mesh/MeshInfo.hpp                             MeshInfo::fromGmshFile, meshCuboid-like generators
//   primitivestorage/SetupPrimitiveStorage.hpp    SetupPrimitiveStorage, setMeshBoundaryFlagsOnBoundary
//   p1functionspace/VertexDoFFunction.hpp         P1Function: interpolate/assign/add/dotGlobal/...
//   operators/Operator.hpp:37-167                 Operator::apply
//   p1functionspace/P1Operator.hpp:192-447        smooth_jac / smooth_sor / smooth_gs
//   solvers/*.hpp                                 Solver, CGSolver, smoothers, GeometricMultigridSolver
//   gridtransferoperators/P1toP1Linear*.hpp       restriction / prolongation
#pragma once

#include <array>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../hyteg_b200.h"

namespace hyteg {

using real_t = double;
using uint_t = std::size_t;
using idx_t  = std::int64_t;

enum UpdateType
{
   Replace = 0,
   Add     = 1
};

enum DoFType : std::size_t
{
   None              = 0,
   All               = 1 + 2 + 4 + 8,
   Boundary          = 2 + 4 + 8,
   Inner             = 1,
   DirichletBoundary = 2,
   NeumannBoundary   = 4,
   FreeslipBoundary  = 8
};

inline DoFType operator|( DoFType a, DoFType b )
{
   return DoFType( static_cast< std::size_t >( a ) | static_cast< std::size_t >( b ) );
}
inline DoFType operator&( DoFType a, DoFType b )
{
   return DoFType( static_cast< std::size_t >( a ) & static_cast< std::size_t >( b ) );
}
inline DoFType operator^( DoFType a, DoFType b )
{
   return DoFType( static_cast< std::size_t >( a ) ^ static_cast< std::size_t >( b ) );
}
inline bool testFlag( DoFType a, DoFType b )
{
   return ( a & b ) != 0;
}

enum class CycleType
{
   VCYCLE,
   WCYCLE
};

// errors are fatal, as in the reference (WALBERLA_ABORT / std::runtime_error)
inline void hbCheck( int rc )
{
   if ( rc != 0 )
      throw std::runtime_error( hb_last_error() );
}
template < typename T >
inline T* hbCheckPtr( T* p )
{
   if ( p == nullptr )
      throw std::runtime_error( hb_last_error() );
   return p;
}

struct Point3D
{
   std::array< real_t, 3 > v{ 0, 0, 0 };
   Point3D() = default;
   Point3D( real_t x, real_t y, real_t z )
   : v{ x, y, z }
   {}
   real_t  operator[]( uint_t i ) const { return v[i]; }
   real_t& operator[]( uint_t i ) { return v[i]; }
   real_t  x() const { return v[0]; }
   real_t  y() const { return v[1]; }
   real_t  z() const { return v[2]; }
   real_t  norm() const { return std::sqrt( v[0] * v[0] + v[1] * v[1] + v[2] * v[2] ); }
};

// ---------------------------------------------------------------------------------------------
// MeshInfo: tetrahedral macro meshes (only what the path needs as input)
// ---------------------------------------------------------------------------------------------
class MeshInfo
{
 public:
   std::vector< real_t > vertices; // xyz triples
   std::vector< int >    cells;    // 4 vertex ids per tetrahedron

   /// Gmsh 2.2 ASCII reader for tetrahedral meshes (element type 4), cf. MeshInfo::fromGmshFile
   static MeshInfo fromGmshFile( const std::string& fileName )
   {
      std::ifstream in( fileName );
      if ( !in )
         throw std::runtime_error( "MeshInfo: cannot open " + fileName );
      MeshInfo    m;
      std::string tok;
      std::vector< long > ids;
      while ( in >> tok )
      {
         if ( tok == "$Nodes" )
         {
            long n;
            in >> n;
            for ( long i = 0; i < n; i++ )
            {
               long   id;
               real_t x, y, z;
               in >> id >> x >> y >> z;
               ids.push_back( id );
               m.vertices.insert( m.vertices.end(), { x, y, z } );
            }
         }
         else if ( tok == "$Elements" )
         {
            long n;
            in >> n;
            std::string line;
            std::getline( in, line );
            for ( long i = 0; i < n; i++ )
            {
               std::getline( in, line );
               std::istringstream ls( line );
               long               id, type, ntags;
               ls >> id >> type >> ntags;
               for ( long t = 0; t < ntags; t++ )
               {
                  long tag;
                  ls >> tag;
               }
               if ( type != 4 )
                  continue;
               for ( int k = 0; k < 4; k++ )
               {
                  long v;
                  ls >> v;
                  int idx = -1;
                  for ( size_t j = 0; j < ids.size(); j++ )
                     if ( ids[j] == v )
                        idx = (int) j;
                  if ( idx < 0 )
                     throw std::runtime_error( "MeshInfo: unknown node id in " + fileName );
                  m.cells.push_back( idx );
               }
            }
         }
      }
      if ( m.cells.empty() )
         throw std::runtime_error( "MeshInfo: no tetrahedra in " + fileName );
      return m;
   }

   /// the unit cube split into six tetrahedra: same vertices and cells as data/meshes/3D/cube_6el.msh
   static MeshInfo meshCube6el()
   {
      MeshInfo m;
      m.vertices = { 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 0, 0, 1, 1, 0, 1, 1, 1, 1, 0, 1, 1 };
      m.cells    = { 0, 1, 3, 4, 1, 3, 4, 5, 1, 2, 3, 5, 4, 5, 3, 7, 5, 6, 3, 7, 5, 2, 3, 6 };
      return m;
   }
};

class SetupPrimitiveStorage
{
 public:
   explicit SetupPrimitiveStorage( const MeshInfo& meshInfo, uint_t numberOfProcesses = 1 )
   : meshInfo_( meshInfo )
   , nProcs_( numberOfProcesses )
   {}
   void setMeshBoundaryFlagsOnBoundary( uint_t flagOnBoundary, uint_t flagInner, bool highestDimensionAlwaysInner )
   {
      haveFlags_ = true;
      fb_        = (int) flagOnBoundary;
      fi_        = (int) flagInner;
      hd_        = highestDimensionAlwaysInner;
   }
   const MeshInfo& meshInfo() const { return meshInfo_; }

 private:
   friend class PrimitiveStorage;
   MeshInfo meshInfo_;
   uint_t   nProcs_;
   bool     haveFlags_ = false;
   int      fb_ = 1, fi_ = 0;
   bool     hd_ = true;
};

class PrimitiveStorage
{
 public:
   explicit PrimitiveStorage( const SetupPrimitiveStorage& setup, uint_t rank = 0 )
   {
      const MeshInfo& m = setup.meshInfo_;
      st_ = hbCheckPtr( hb_storage_create( (int) ( m.vertices.size() / 3 ), m.vertices.data(), (int) ( m.cells.size() / 4 ), m.cells.data(), (int) rank, (int) setup.nProcs_ ) );
      if ( setup.haveFlags_ )
         hbCheck( hb_storage_set_mesh_boundary_flags_on_boundary( st_, setup.fb_, setup.fi_, setup.hd_ ? 1 : 0 ) );
   }
   /// non-owning view of an existing C-ABI handle
   explicit PrimitiveStorage( hb_storage* borrowed )
   : st_( borrowed )
   , owned_( false )
   {}
   ~PrimitiveStorage()
   {
      if ( owned_ )
         hb_storage_destroy( st_ );
   }
   PrimitiveStorage( const PrimitiveStorage& )            = delete;
   PrimitiveStorage& operator=( const PrimitiveStorage& ) = delete;

   bool        hasGlobalCells() const { return true; }
   uint_t      getNumberOfLocalCells() const { return (uint_t) hb_storage_num_local_cells( st_ ); }
   hb_storage* handle() const { return st_; }

 private:
   hb_storage* st_;
   bool        owned_ = true;
};

inline uint_t numberOfGlobalDoFsP1( const PrimitiveStorage& storage, uint_t level )
{
   return (uint_t) hb_storage_num_global_dofs( storage.handle(), (int) level );
}

// ---------------------------------------------------------------------------------------------
// P1Function
// ---------------------------------------------------------------------------------------------
template < typename ValueType = real_t >
class P1Function
{
   static_assert( std::is_same< ValueType, double >::value, "the B200 path is FP64, as the reference's 3D kernels (P1ConstantOperator.cpp:346-349)" );

 public:
   using valueType = ValueType;

   P1Function( const std::string& name, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : name_( name )
   , storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , f_( hbCheckPtr( hb_p1function_create( storage->handle(), (int) minLevel, (int) maxLevel ) ) )
   {}
   /// non-owning view of an existing C-ABI handle
   P1Function( hb_function* borrowed, const std::shared_ptr< PrimitiveStorage >& storage, uint_t minLevel, uint_t maxLevel )
   : name_( "borrowed" )
   , storage_( storage )
   , minLevel_( minLevel )
   , maxLevel_( maxLevel )
   , f_( borrowed )
   , owned_( false )
   {}
   ~P1Function()
   {
      if ( owned_ )
         hb_p1function_destroy( f_ );
   }
   P1Function( const P1Function& )            = delete;
   P1Function& operator=( const P1Function& ) = delete;

   const std::string&                  getFunctionName() const { return name_; }
   std::shared_ptr< PrimitiveStorage > getStorage() const { return storage_; }
   uint_t                              getMinLevel() const { return minLevel_; }
   uint_t                              getMaxLevel() const { return maxLevel_; }
   hb_function*                        handle() const { return f_; }

   void interpolate( ValueType constant, uint_t level, DoFType flag = All ) const
   {
      hbCheck( hb_p1function_interpolate_constant( f_, constant, (int) level, (int) flag ) );
   }
   void interpolate( const std::function< ValueType( const Point3D& ) >& expr, uint_t level, DoFType flag = All ) const
   {
      const size_t n = (size_t) hb_storage_num_local_cells( storage_->handle() ) * (size_t) hb_cell_array_size( (int) level );
      std::vector< double > xyz( 3 * n ), val( n );
      hbCheck( hb_storage_micro_coordinates( storage_->handle(), (int) level, xyz.data() ) );
      for ( size_t i = 0; i < n; i++ )
         val[i] = expr( Point3D( xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2] ) );
      hbCheck( hb_p1function_upload( f_, (int) level, val.data(), (int) flag ) );
   }
   void setToZero( uint_t level ) const { interpolate( ValueType( 0 ), level, All ); }

   void assign( const std::vector< ValueType >& scalars, const std::vector< std::reference_wrapper< const P1Function > >& functions, uint_t level, DoFType flag = All ) const
   {
      std::vector< hb_function* > h;
      for ( const P1Function& f : functions )
         h.push_back( f.f_ );
      hbCheck( hb_p1function_assign( f_, (int) h.size(), scalars.data(), h.data(), (int) level, (int) flag ) );
   }
   void add( const std::vector< ValueType >& scalars, const std::vector< std::reference_wrapper< const P1Function > >& functions, uint_t level, DoFType flag = All ) const
   {
      std::vector< hb_function* > h;
      for ( const P1Function& f : functions )
         h.push_back( f.f_ );
      hbCheck( hb_p1function_add( f_, (int) h.size(), scalars.data(), h.data(), (int) level, (int) flag ) );
   }
   void add( ValueType scalar, uint_t level, DoFType flag = All ) const { hbCheck( hb_p1function_add_scalar( f_, scalar, (int) level, (int) flag ) ); }
   void multElementwise( const std::vector< std::reference_wrapper< const P1Function > >& functions, uint_t level, DoFType flag = All ) const
   {
      std::vector< hb_function* > h;
      for ( const P1Function& f : functions )
         h.push_back( f.f_ );
      hbCheck( hb_p1function_mult_elementwise( f_, (int) h.size(), h.data(), (int) level, (int) flag ) );
   }
   ValueType dotGlobal( const P1Function& rhs, uint_t level, DoFType flag = All ) const
   {
      double v;
      hbCheck( hb_p1function_dot( f_, rhs.f_, (int) level, (int) flag, &v ) );
      return v;
   }
   ValueType sumGlobal( uint_t level, DoFType flag = All ) const
   {
      double v;
      hbCheck( hb_p1function_sum( f_, (int) level, (int) flag, &v ) );
      return v;
   }
   ValueType getMaxDoFMagnitude( uint_t level, DoFType flag = All ) const
   {
      double v;
      hbCheck( hb_p1function_max_magnitude( f_, (int) level, (int) flag, &v ) );
      return v;
   }
   void copyBoundaryConditionFromFunction( const P1Function& other ) const { hbCheck( hb_p1function_copy_boundary_condition( f_, other.f_ ) ); }
   void setBoundaryCondition( const std::vector< DoFType >& typeOfMeshBoundaryFlag ) const
   {
      std::vector< int > t;
      for ( DoFType d : typeOfMeshBoundaryFlag )
         t.push_back( (int) d );
      hbCheck( hb_p1function_set_boundary_condition( f_, t.data(), (int) t.size() ) );
   }

 private:
   std::string                         name_;
   std::shared_ptr< PrimitiveStorage > storage_;
   uint_t                              minLevel_, maxLevel_;
   hb_function*                        f_;
   bool                                owned_ = true;
};

template < typename F >
inline void copyBCs( const F& src, const F& dst )
{
   dst.copyBoundaryConditionFromFunction( src );
}

} // namespace hyteg

#include "hyteg_solvers.hpp"
