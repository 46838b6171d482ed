// HyTeG-side bridge: hyteg::P1ConstantLaplaceOperatorB200, a drop-in for hyteg::P1ConstantLaplaceOperator whose apply / smooth_* run in
// libhyteg_b200.so.  This file is compiled AGAINST THE REFERENCE'S OWN HEADERS (it is what a HyTeG maintainer would add to the tree); it
// derives from the reference's Operator< P1Function, P1Function > (src/hyteg/operators/Operator.hpp:37) and the Smoothables of
// src/hyteg/solvers/Smoothables.hpp:33-123, so that the reference's GaussSeidelSmoother, SORSmoother, WeightedJacobiSmoother,
// ChebyshevSmoother, CGSolver and GeometricMultigridSolver templates accept it unchanged.
//
// Data crosses the boundary one macro-cell array at a time (hb_p1function_upload_cell / download_cell): the library's per-cell layout IS
// HyTeG's macro-cell layout (indexing/MacroCellIndexing.hpp:44-52), so a transfer is a memcpy of the reference's own buffer after the
// usual Vertex -> Edge -> Face -> Cell communication has completed the cell arrays.  On the way back the lower-dimensional primitives'
// own data is read off a neighbouring cell's array with the reference's own iterators (the reverse of
// VertexDoFPackInfo::communicateLocal{Vertex,Edge,Face}ToCell, p1functionspace/VertexDoFPackInfo.hpp:438-505, 710-813).
//
// This "copy per call" mode exists to run the reference's unmodified drivers and tests on the device path; a production integration keeps
// the functions device-resident and crosses the boundary once per solve.
#pragma once

#include <algorithm>
#include <map>
#include <memory>
#include <set>
#include <vector>

#include "core/Abort.h"
#include "core/DataTypes.h"

#include "hyteg/Levelinfo.hpp"
#include "hyteg/operators/Operator.hpp"
#include "hyteg/p1functionspace/P1Function.hpp"
#include "hyteg/p1functionspace/VertexDoFIndexing.hpp"
#include "hyteg/p1functionspace/VertexDoFMacroCell.hpp"
#include "hyteg/p1functionspace/VertexDoFMacroFace.hpp"
#include "hyteg/primitivestorage/PrimitiveStorage.hpp"
#include "hyteg/solvers/Smoothables.hpp"

#include "constant_stencil_operator/P1ConstantOperator.hpp" // 2D storages stay on the reference's own operator

#include "hyteg_b200.h"

namespace hyteg {

#define HB_BRIDGE_CHECK( call )                                       \
   do                                                                 \
   {                                                                  \
      if ( ( call ) != 0 )                                            \
         WALBERLA_ABORT( "hyteg_b200: " << hb_last_error() );         \
   } while ( 0 )

/// the device twin of one PrimitiveStorage: macro mesh, boundary flags and the P1Function mirrors, shared by all bridge operators on it
class B200StorageMirror
{
 public:
   explicit B200StorageMirror( const std::shared_ptr< PrimitiveStorage >& storage )
   : storage_( storage )
   {
      WALBERLA_CHECK( storage->hasGlobalCells(), "hyteg_b200 covers the 3D (macro-cell) path" );
      WALBERLA_CHECK_EQUAL( walberla::mpi::MPIManager::instance()->numProcesses(), 1, "the copy-per-call bridge is a single-rank tool" );

      // macro vertices in PrimitiveID order -> 0..nV-1
      std::vector< PrimitiveID > vids = storage->getVertexIDs();
      std::sort( vids.begin(), vids.end() );
      std::vector< double > xyz;
      for ( const auto& id : vids )
      {
         vertexIndex_[id] = int( vertexIndex_.size() );
         const Point3D c  = storage->getVertex( id )->getCoordinates();
         xyz.insert( xyz.end(), { double( c[0] ), double( c[1] ), double( c[2] ) } );
      }
      // macro-cells: the four vertices in the cell's LOCAL order, which fixes the x / y / z axes of its array
      cellIDs_ = storage->getCellIDs();
      std::sort( cellIDs_.begin(), cellIDs_.end() );
      std::vector< int > cells;
      for ( const auto& id : cellIDs_ )
         for ( const auto& v : storage->getCell( id )->neighborVertices() )
            cells.push_back( vertexIndex_.at( v ) );
      st_ = hb_storage_create( int( vids.size() ), xyz.data(), int( cellIDs_.size() ), cells.data(), 0, 1 );
      if ( !st_ )
         WALBERLA_ABORT( "hyteg_b200: " << hb_last_error() );

      // Primitive::getMeshBoundaryFlag() of every primitive, looked up through its vertex set
      std::map< std::vector< int >, int > flagOf;
      auto                                 note = [&]( const Primitive& p, const std::vector< PrimitiveID >& verts ) {
         std::vector< int > key;
         for ( const auto& v : verts )
            key.push_back( vertexIndex_.at( v ) );
         std::sort( key.begin(), key.end() );
         flagOf[key] = int( p.getMeshBoundaryFlag() );
         maxFlag_    = std::max( maxFlag_, int( p.getMeshBoundaryFlag() ) );
      };
      for ( const auto& it : storage->getVertices() )
         note( *it.second, { it.first } );
      for ( const auto& it : storage->getEdges() )
         note( *it.second, it.second->neighborVertices() );
      for ( const auto& it : storage->getFaces() )
         note( *it.second, it.second->neighborVertices() );
      for ( const auto& it : storage->getCells() )
         note( *it.second, it.second->neighborVertices() );
      int counts[4];
      HB_BRIDGE_CHECK( hb_storage_counts( st_, counts ) );
      for ( int dim = 0; dim < 4; dim++ )
      {
         std::vector< int > verts( size_t( counts[dim] ) * size_t( dim + 1 ) ), flags( size_t( counts[dim] ) );
         HB_BRIDGE_CHECK( hb_storage_primitive_vertices( st_, dim, verts.data() ) );
         for ( int i = 0; i < counts[dim]; i++ )
         {
            std::vector< int > key( verts.begin() + i * ( dim + 1 ), verts.begin() + ( i + 1 ) * ( dim + 1 ) );
            std::sort( key.begin(), key.end() );
            flags[size_t( i )] = flagOf.at( key );
         }
         HB_BRIDGE_CHECK( hb_storage_set_mesh_boundary_flags( st_, dim, flags.data() ) );
      }
   }
   ~B200StorageMirror()
   {
      for ( auto& m : mirrors_ )
         hb_p1function_destroy( m.second );
      hb_storage_destroy( st_ );
   }
   B200StorageMirror( const B200StorageMirror& )            = delete;
   B200StorageMirror& operator=( const B200StorageMirror& ) = delete;

   hb_storage* handle() const { return st_; }

   /// one mirror per (storage, function): created on first use with the function's level range and BoundaryCondition table
   hb_function* mirror( const P1Function< real_t >& f )
   {
      const uint_t key = uint_t( f.getCellDataID() );
      auto         it  = mirrors_.find( key );
      if ( it != mirrors_.end() )
         return it->second;
      hb_function* m = hb_p1function_create( st_, int( f.getMinLevel() ), int( f.getMaxLevel() ) );
      if ( !m )
         WALBERLA_ABORT( "hyteg_b200: " << hb_last_error() );
      std::vector< int >      table;
      const BoundaryCondition bc = f.getBoundaryCondition();
      for ( int flag = 0; flag <= maxFlag_; flag++ )
         table.push_back( int( bc.getBoundaryType( uint_c( flag ) ) ) );
      HB_BRIDGE_CHECK( hb_p1function_set_boundary_condition( m, table.data(), int( table.size() ) ) );
      mirrors_[key] = m;
      return m;
   }

   /// host -> device: complete the macro-cell arrays the way every reference kernel does before it runs, then copy them
   hb_function* upload( const P1Function< real_t >& f, uint_t level )
   {
      hb_function* m = mirror( f );
      f.communicate< Vertex, Edge >( level );
      f.communicate< Edge, Face >( level );
      f.communicate< Face, Cell >( level );
      for ( size_t c = 0; c < cellIDs_.size(); c++ )
      {
         const Cell* cell = storage_->getCell( cellIDs_[c] );
         HB_BRIDGE_CHECK( hb_p1function_upload_cell( m, int( level ), int( c ), cell->getData( f.getCellDataID() )->getPointer( level ) ) );
      }
      return m;
   }

   /// device -> host: the macro-cell arrays, then every lower-dimensional primitive's own data from one of its cells
   void download( const P1Function< real_t >& f, uint_t level )
   {
      hb_function* m = mirror( f );
      for ( size_t c = 0; c < cellIDs_.size(); c++ )
      {
         Cell* cell = storage_->getCell( cellIDs_[c] );
         HB_BRIDGE_CHECK( hb_p1function_download_cell( m, int( level ), int( c ), cell->getData( f.getCellDataID() )->getPointer( level ) ) );
      }
      const uint_t n = levelinfo::num_microvertices_per_edge( level );
      for ( const auto& it : storage_->getFaces() )
      {
         Face&         face = *it.second;
         const Cell&   cell = *storage_->getCell( face.neighborCells()[0] );
         const real_t* cd   = cell.getData( f.getCellDataID() )->getPointer( level );
         real_t*       fd   = face.getData( f.getFaceDataID() )->getPointer( level );
         const uint_t  lf   = cell.getLocalFaceID( face.getID() );
         const auto&   map  = cell.getFaceLocalVertexToCellLocalVertexMaps().at( lf );
         auto          ci   = vertexdof::macrocell::BoundaryIterator( level, map.at( 0 ), map.at( 1 ), map.at( 2 ) );
         for ( const auto& fi : vertexdof::macroface::Iterator( level ) )
         {
            fd[vertexdof::macroface::index( level, fi.x(), fi.y() )] = cd[vertexdof::macrocell::index( level, ci->x(), ci->y(), ci->z() )];
            ci++;
         }
      }
      for ( const auto& it : storage_->getEdges() )
      {
         Edge&              edge = *it.second;
         const Cell&        cell = *storage_->getCell( edge.neighborCells()[0] );
         const real_t*      cd   = cell.getData( f.getCellDataID() )->getPointer( level );
         real_t*            ed   = edge.getData( f.getEdgeDataID() )->getPointer( level );
         const uint_t       le   = cell.getLocalEdgeID( edge.getID() );
         const auto&        map  = cell.getEdgeLocalVertexToCellLocalVertexMaps().at( le );
         std::set< uint_t > rest = { 0, 1, 2, 3 };
         rest.erase( map.at( 0 ) );
         rest.erase( map.at( 1 ) );
         auto ci = vertexdof::macrocell::BoundaryIterator( level, map.at( 0 ), map.at( 1 ), *rest.begin(), 0 );
         for ( uint_t i = 0; i < n; i++ )
         {
            ed[i] = cd[vertexdof::macrocell::index( level, ci->x(), ci->y(), ci->z() )];
            ci++;
         }
      }
      for ( const auto& it : storage_->getVertices() )
      {
         Vertex&       vertex = *it.second;
         const Cell&   cell   = *storage_->getCell( vertex.neighborCells()[0] );
         const real_t* cd     = cell.getData( f.getCellDataID() )->getPointer( level );
         const uint_t  lv     = cell.getLocalVertexID( vertex.getID() );
         const idx_t   e      = idx_t( n ) - 1;
         vertex.getData( f.getVertexDataID() )->getPointer( level )[0] =
             cd[vertexdof::macrocell::index( level, lv == 1 ? e : 0, lv == 2 ? e : 0, lv == 3 ? e : 0 )];
      }
   }

   /// one twin per PrimitiveStorage, shared by every bridge operator / function on it
   static std::shared_ptr< B200StorageMirror > of( const std::shared_ptr< PrimitiveStorage >& storage )
   {
      static std::map< const PrimitiveStorage*, std::weak_ptr< B200StorageMirror > > table;
      auto                                                                           sp = table[storage.get()].lock();
      if ( !sp )
      {
         sp                   = std::make_shared< B200StorageMirror >( storage );
         table[storage.get()] = sp;
      }
      return sp;
   }

 private:
   std::shared_ptr< PrimitiveStorage >  storage_;
   hb_storage*                          st_ = nullptr;
   std::map< PrimitiveID, int >         vertexIndex_;
   std::vector< PrimitiveID >           cellIDs_;
   std::map< uint_t, hb_function* >     mirrors_;
   int                                  maxFlag_ = 0;
};

class P1ConstantLaplaceOperatorB200 : public Operator< P1Function< real_t >, P1Function< real_t > >,
                                      public GSSmoothable< P1Function< real_t > >,
                                      public GSBackwardsSmoothable< P1Function< real_t > >,
                                      public SORSmoothable< P1Function< real_t > >,
                                      public SORBackwardsSmoothable< P1Function< real_t > >,
                                      public WeightedJacobiSmoothable< P1Function< real_t > >,
                                      public ConstantJacobiSmoothable< P1Function< real_t > >,
                                      public OperatorWithInverseDiagonal< P1Function< real_t > >
{
 public:
   P1ConstantLaplaceOperatorB200( const std::shared_ptr< PrimitiveStorage >& storage, size_t minLevel, size_t maxLevel )
   : Operator( storage, minLevel, maxLevel )
   {
      if ( !storage->hasGlobalCells() )
      {
         // the library covers the 3D (macro-cell) path; a 2D macro mesh keeps the reference's operator, so that drivers which loop over
         // 2D and 3D meshes run unchanged
         ref2D_ = std::make_shared< P1ConstantLaplaceOperator >( storage, minLevel, maxLevel );
         return;
      }
      dev_ = B200StorageMirror::of( storage );
      op_  = hb_p1_constant_laplace_create( dev_->handle(), int( minLevel ), int( maxLevel ) );
      if ( !op_ )
         WALBERLA_ABORT( "hyteg_b200: " << hb_last_error() );
   }
   ~P1ConstantLaplaceOperatorB200() override
   {
      if ( op_ )
         hb_operator_destroy( op_ );
   }
   P1ConstantLaplaceOperatorB200( const P1ConstantLaplaceOperatorB200& )            = delete;
   P1ConstantLaplaceOperatorB200& operator=( const P1ConstantLaplaceOperatorB200& ) = delete;

   /// P1ConstantOperator::apply (constant_stencil_operator/P1ConstantOperator.cpp:380-527)
   void apply( const P1Function< real_t >& src,
               const P1Function< real_t >& dst,
               size_t                      level,
               DoFType                     flag,
               UpdateType                  updateType = Replace ) const override
   {
      if ( ref2D_ )
         return ref2D_->apply( src, dst, level, flag, updateType );
      hb_function* s = dev_->upload( src, level );
      hb_function* d = dev_->upload( dst, level ); // Add needs dst; Replace keeps the DoFs outside `flag`
      HB_BRIDGE_CHECK( hb_operator_apply( op_, s, d, int( level ), int( flag ), updateType == Replace ? 0 : 1 ) );
      dev_->download( dst, level );
   }

   /// P1ConstantOperator::smooth_sor (P1ConstantOperator.cpp:596-790); the library's sweep is the multicolour ordering of DESIGN.md section 3
   void smooth_sor( const P1Function< real_t >& dst, const P1Function< real_t >& rhs, real_t relax, size_t level, DoFType flag ) const override
   {
      sor( dst, rhs, relax, level, flag, 0 );
   }
   void smooth_sor_backwards( const P1Function< real_t >& dst, const P1Function< real_t >& rhs, real_t relax, size_t level, DoFType flag ) const override
   {
      sor( dst, rhs, relax, level, flag, 1 );
   }
   void smooth_gs( const P1Function< real_t >& dst, const P1Function< real_t >& rhs, size_t level, DoFType flag ) const override
   {
      sor( dst, rhs, real_c( 1 ), level, flag, 0 );
   }
   void smooth_gs_backwards( const P1Function< real_t >& dst, const P1Function< real_t >& rhs, size_t level, DoFType flag ) const override
   {
      sor( dst, rhs, real_c( 1 ), level, flag, 1 );
   }
   /// P1ConstantOperator::smooth_jac (P1ConstantOperator.cpp:897-960): dst = (1 - relax) src + relax D^-1 ( rhs - (A - D) src )
   void smooth_jac( const P1Function< real_t >& dst,
                    const P1Function< real_t >& rhs,
                    const P1Function< real_t >& src,
                    real_t                      relax,
                    size_t                      level,
                    DoFType                     flag ) const override
   {
      if ( ref2D_ )
         return ref2D_->smooth_jac( dst, rhs, src, relax, level, flag );
      hb_function* d = dev_->upload( dst, level );
      hb_function* r = dev_->upload( rhs, level );
      hb_function* s = dev_->upload( src, level );
      HB_BRIDGE_CHECK( hb_operator_smooth_jac( op_, d, r, s, double( relax ), int( level ), int( flag ) ) );
      dev_->download( dst, level );
   }
   void smooth_jac( const P1Function< real_t >& dst, const P1Function< real_t >& rhs, const P1Function< real_t >& src, size_t level, DoFType flag )
       const override
   {
      smooth_jac( dst, rhs, src, real_c( 1 ), level, flag );
   }

   /// P1Operator::computeInverseDiagonalOperatorValues (p1functionspace/P1Operator.hpp:178-200): all levels, from the device stencils
   void computeInverseDiagonalOperatorValues() override
   {
      if ( ref2D_ )
      {
         ref2D_->computeInverseDiagonalOperatorValues();
         inverseDiagonal_ = ref2D_->getInverseDiagonalValues();
         return;
      }
      if ( !inverseDiagonal_ )
         inverseDiagonal_ = std::make_shared< P1Function< real_t > >( "inverse diagonal (b200)", storage_, minLevel_, maxLevel_ );
      for ( uint_t level = minLevel_; level <= maxLevel_; level++ )
      {
         hb_function* d = dev_->mirror( *inverseDiagonal_ );
         HB_BRIDGE_CHECK( hb_operator_diagonal( op_, d, int( level ), 1 ) );
         dev_->download( *inverseDiagonal_, level );
      }
   }
   std::shared_ptr< P1Function< real_t > > getInverseDiagonalValues() const override
   {
      WALBERLA_CHECK( inverseDiagonal_ != nullptr,
                      "Inverse diagonal values have not been assembled, call computeInverseDiagonalOperatorValues() to set up this function." )
      return inverseDiagonal_;
   }

 private:
   void sor( const P1Function< real_t >& dst, const P1Function< real_t >& rhs, real_t relax, size_t level, DoFType flag, int backwards ) const
   {
      if ( ref2D_ )
      {
         if ( backwards )
            ref2D_->smooth_sor_backwards( dst, rhs, relax, level, flag );
         else
            ref2D_->smooth_sor( dst, rhs, relax, level, flag );
         return;
      }
      hb_function* d = dev_->upload( dst, level );
      hb_function* r = dev_->upload( rhs, level );
      HB_BRIDGE_CHECK( hb_operator_smooth_sor( op_, d, r, double( relax ), int( level ), int( flag ), backwards ) );
      dev_->download( dst, level );
   }

   std::shared_ptr< B200StorageMirror >         dev_;
   hb_operator*                                 op_ = nullptr;
   std::shared_ptr< P1ConstantLaplaceOperator > ref2D_; // set instead of dev_ / op_ on 2D storages
   std::shared_ptr< P1Function< real_t > > inverseDiagonal_;
};

} // namespace hyteg
