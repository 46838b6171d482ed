#include "topology.hpp"

#include <algorithm>
#include <set>
#include <stdexcept>

namespace hb {

const int kLocalEdgeV[6][2] = { { 0, 1 }, { 0, 2 }, { 1, 2 }, { 0, 3 }, { 1, 3 }, { 2, 3 } };
const int kLocalFaceV[4][3] = { { 0, 1, 2 }, { 0, 1, 3 }, { 0, 2, 3 }, { 1, 2, 3 } };

//                              0   1   2   3   4   5   6   7   8   9  10  11  12  13  14  15
const int kFaceOfMask[16]   = { -1, 2, 1, -1, 0, -1, -1, -1, 3, -1, -1, -1, -1, -1, -1, -1 };
const int kEdgeOfMask[16]   = { -1, -1, -1, 3, -1, 1, 0, -1, -1, 5, 4, -1, 2, -1, -1, -1 };
const int kVertexOfMask[16] = { -1, -1, -1, -1, -1, -1, -1, 0, -1, -1, -1, 3, -1, 2, 1, -1 };

void Topology::build( int nVertices, const double* coords, int nCells, const int* cells )
{
   if ( nVertices <= 0 || nCells <= 0 )
      throw std::runtime_error( "hyteg_b200: empty macro mesh" );
   nV = nVertices;
   nC = nCells;
   xyz.assign( coords, coords + 3 * nV );
   cellV.resize( nC );
   for ( int c = 0; c < nC; c++ )
      for ( int j = 0; j < 4; j++ )
      {
         int v = cells[4 * c + j];
         if ( v < 0 || v >= nV )
            throw std::runtime_error( "hyteg_b200: cell vertex id out of range" );
         cellV[c][j] = v;
      }

   std::map< std::array< int, 2 >, int > edgeMap;
   std::map< std::array< int, 3 >, int > faceMap;
   cellE.resize( nC );
   cellF.resize( nC );
   vCells.assign( nV, {} );
   for ( int c = 0; c < nC; c++ )
   {
      for ( int j = 0; j < 4; j++ )
         vCells[cellV[c][j]].push_back( c );
      for ( int e = 0; e < 6; e++ )
      {
         std::array< int, 2 > k = { cellV[c][kLocalEdgeV[e][0]], cellV[c][kLocalEdgeV[e][1]] };
         std::sort( k.begin(), k.end() );
         auto it = edgeMap.find( k );
         if ( it == edgeMap.end() )
         {
            it = edgeMap.emplace( k, (int) edgeV.size() ).first;
            edgeV.push_back( k );
         }
         cellE[c][e] = it->second;
      }
      for ( int f = 0; f < 4; f++ )
      {
         std::array< int, 3 > k = { cellV[c][kLocalFaceV[f][0]], cellV[c][kLocalFaceV[f][1]], cellV[c][kLocalFaceV[f][2]] };
         std::sort( k.begin(), k.end() );
         auto it = faceMap.find( k );
         if ( it == faceMap.end() )
         {
            it = faceMap.emplace( k, (int) faceV.size() ).first;
            faceV.push_back( k );
         }
         cellF[c][f] = it->second;
      }
   }
   nE = (int) edgeV.size();
   nF = (int) faceV.size();

   vNumCells.assign( nV, 0 );
   eNumCells.assign( nE, 0 );
   fNumCells.assign( nF, 0 );
   vMinCell.assign( nV, nC );
   eMinCell.assign( nE, nC );
   fMinCell.assign( nF, nC );
   for ( int c = 0; c < nC; c++ )
   {
      for ( int j = 0; j < 4; j++ )
      {
         vNumCells[cellV[c][j]]++;
         vMinCell[cellV[c][j]] = std::min( vMinCell[cellV[c][j]], c );
      }
      for ( int e = 0; e < 6; e++ )
      {
         eNumCells[cellE[c][e]]++;
         eMinCell[cellE[c][e]] = std::min( eMinCell[cellE[c][e]], c );
      }
      for ( int f = 0; f < 4; f++ )
      {
         fNumCells[cellF[c][f]]++;
         fMinCell[cellF[c][f]] = std::min( fMinCell[cellF[c][f]], c );
      }
   }
   for ( int f = 0; f < nF; f++ )
      if ( fNumCells[f] > 2 )
         throw std::runtime_error( "hyteg_b200: macro face with more than two neighbour cells" );

   // boundary detection, reference SetupPrimitiveStorage::onBoundary (SetupPrimitiveStorage.cpp:1092-)
   fOnBnd.assign( nF, 0 );
   eOnBnd.assign( nE, 0 );
   vOnBnd.assign( nV, 0 );
   cOnBnd.assign( nC, 0 );
   for ( int f = 0; f < nF; f++ )
      fOnBnd[f] = fNumCells[f] == 1;
   for ( int c = 0; c < nC; c++ )
      for ( int f = 0; f < 4; f++ )
      {
         int gf = cellF[c][f];
         if ( !fOnBnd[gf] )
            continue;
         cOnBnd[c] = 1;
         for ( int a = 0; a < 3; a++ )
         {
            int la = kLocalFaceV[f][a];
            vOnBnd[cellV[c][la]] = 1;
            for ( int b = a + 1; b < 3; b++ )
            {
               int lb = kLocalFaceV[f][b];
               for ( int e = 0; e < 6; e++ )
                  if ( kLocalEdgeV[e][0] == la && kLocalEdgeV[e][1] == lb )
                     eOnBnd[cellE[c][e]] = 1;
            }
         }
      }
   vFlag.assign( nV, 0 );
   eFlag.assign( nE, 0 );
   fFlag.assign( nF, 0 );
   cFlag.assign( nC, 0 );
   cellRank.assign( nC, 0 );
}

void Topology::setMeshBoundaryFlagsOnBoundary( int flagOnBoundary, int flagInner, bool highestDimAlwaysInner )
{
   for ( int i = 0; i < nV; i++ )
      vFlag[i] = vOnBnd[i] ? flagOnBoundary : flagInner;
   for ( int i = 0; i < nE; i++ )
      eFlag[i] = eOnBnd[i] ? flagOnBoundary : flagInner;
   for ( int i = 0; i < nF; i++ )
      fFlag[i] = fOnBnd[i] ? flagOnBoundary : flagInner;
   for ( int i = 0; i < nC; i++ )
      cFlag[i] = ( cOnBnd[i] && !highestDimAlwaysInner ) ? flagOnBoundary : flagInner;
}

void Topology::partition( int nRanks )
{
   if ( nRanks < 1 )
      throw std::runtime_error( "hyteg_b200: nRanks < 1" );
   // contiguous blocks of macro-cells per rank
   for ( int c = 0; c < nC; c++ )
      cellRank[c] = (int) ( ( (int64_t) c * nRanks ) / nC );
}

void Topology::regionPrimitive( int cell, int mask, int& dim, int& id ) const
{
   if ( mask == 0 )
   {
      dim = 3;
      id  = cell;
   }
   else if ( kFaceOfMask[mask] >= 0 )
   {
      dim = 2;
      id  = cellF[cell][kFaceOfMask[mask]];
   }
   else if ( kEdgeOfMask[mask] >= 0 )
   {
      dim = 1;
      id  = cellE[cell][kEdgeOfMask[mask]];
   }
   else if ( kVertexOfMask[mask] >= 0 )
   {
      dim = 0;
      id  = cellV[cell][kVertexOfMask[mask]];
   }
   else
   {
      dim = 255;
      id  = -1;
   }
}

// lowest global id among the macro-cells of 'rank' that contain the macro primitive 'region m' of macro-cell g (-1: none)
static int localMinCell( const Topology& topo, int g, int m, int rank )
{
   static const int faceBitOfB[4] = { 8, 1, 2, 4 };
   int              supp[4], ns = 0;
   for ( int j = 0; j < 4; j++ )
      if ( !( m & faceBitOfB[j] ) )
         supp[ns++] = topo.cellV[g][j];
   if ( ns == 0 )
      return -1;
   int best = -1;
   for ( int c : topo.vCells[supp[0]] ) // sorted ascending
   {
      if ( topo.cellRank[c] != rank )
         continue;
      bool all = true;
      for ( int k = 1; k < ns && all; k++ )
      {
         bool has = false;
         for ( int j = 0; j < 4; j++ )
            has = has || topo.cellV[c][j] == supp[k];
         all = has;
      }
      if ( all && ( best < 0 || c < best ) )
         best = c;
   }
   return best;
}

// true if some macro-cell containing the macro primitive 'region m' of macro-cell g belongs to another rank
static bool hasRemoteCell( const Topology& topo, int g, int m, int rank )
{
   static const int faceBitOfB[4] = { 8, 1, 2, 4 };
   int              supp[4], ns = 0;
   for ( int j = 0; j < 4; j++ )
      if ( !( m & faceBitOfB[j] ) )
         supp[ns++] = topo.cellV[g][j];
   if ( ns == 0 )
      return false;
   for ( int c : topo.vCells[supp[0]] )
   {
      if ( topo.cellRank[c] == rank )
         continue;
      bool all = true;
      for ( int k = 1; k < ns && all; k++ )
      {
         bool has = false;
         for ( int j = 0; j < 4; j++ )
            has = has || topo.cellV[c][j] == supp[k];
         all = has;
      }
      if ( all )
         return true;
   }
   return false;
}

uint32_t sharedRegionMasks( const Topology& topo, int g, int rank )
{
   static const int faceBitOfB[4] = { 8, 1, 2, 4 }; // local vertex j is OFF the face with this mask bit
   uint32_t         out           = 0;
   for ( int m = 1; m < HB_NMASK - 1; m++ )
   {
      int dim, id;
      topo.regionPrimitive( g, m, dim, id );
      if ( dim > 2 )
         continue;
      // cells of 'rank' around the first support vertex that contain all support vertices
      int supp[4], ns = 0;
      for ( int j = 0; j < 4; j++ )
         if ( !( m & faceBitOfB[j] ) )
            supp[ns++] = topo.cellV[g][j];
      bool found = false;
      for ( int c : topo.vCells[supp[0]] )
      {
         if ( c == g || topo.cellRank[c] != rank )
            continue;
         bool all = true;
         for ( int k = 1; k < ns && all; k++ )
         {
            bool has = false;
            for ( int j = 0; j < 4; j++ )
               has = has || topo.cellV[c][j] == supp[k];
            all = has;
         }
         if ( all )
         {
            found = true;
            break;
         }
      }
      if ( found )
         out |= 1u << m;
   }
   return out;
}

void LocalView::build( const Topology& topo, int rank_, int nRanks_ )
{
   rank   = rank_;
   nRanks = nRanks_;
   slotGlobal.clear();
   globalToSlot.clear();
   for ( int c = 0; c < topo.nC; c++ )
      if ( topo.cellRank[c] == rank )
      {
         globalToSlot[c] = (int) slotGlobal.size();
         slotGlobal.push_back( c );
      }
   nLocal = (int) slotGlobal.size();

   cells.assign( nLocal, HbCellInfo{} );
   copies.clear();
   regionCopies.clear();
   auto fillRegions = [&topo, this]( HbCellInfo& ci, int c ) {
      ci.globalId = c;
      for ( int j = 0; j < 4; j++ )
         ci.gv[j] = topo.cellV[c][j];
      for ( int m = 0; m < 16; m++ )
      {
         int dim, id;
         topo.regionPrimitive( c, m, dim, id );
         HbRegion& r = ci.region[m];
         r.dim       = (uint8_t) dim;
         switch ( dim )
         {
         case 3:
            r.bflag = (uint8_t) topo.cFlag[c];
            r.owner = 1;
            r.mult  = 1;
            break;
         case 2:
            r.bflag = (uint8_t) topo.fFlag[id];
            r.owner = topo.fMinCell[id] == c;
            r.mult  = (uint8_t) topo.fNumCells[id];
            break;
         case 1:
            r.bflag = (uint8_t) topo.eFlag[id];
            r.owner = topo.eMinCell[id] == c;
            r.mult  = (uint8_t) std::min( 255, topo.eNumCells[id] );
            break;
         case 0:
            r.bflag = (uint8_t) topo.vFlag[id];
            r.owner = topo.vMinCell[id] == c;
            r.mult  = (uint8_t) std::min( 255, topo.vNumCells[id] );
            break;
         default:
            r.bflag = 0;
            r.owner = 0;
            r.mult  = 1;
         }
         // bit 0: owning copy over all ranks (every DoF counted once in dotGlobal); bit 1: owning copy on THIS rank -- that
         // macro-cell evaluates the interface stencil and writes the result to all copies held by the rank
         if ( dim == 3 )
            r.owner = 3;
         else if ( dim <= 2 && localMinCell( topo, c, m, rank ) == c )
            r.owner |= 2;
         // bit 2: the primitive is shared with another rank (its stencil reads mirrored data: needs a current ghost layer)
         if ( dim <= 2 && nRanks > 1 && hasRemoteCell( topo, c, m, rank ) )
            r.owner |= 4;
      }
   };
   for ( int l = 0; l < nLocal; l++ )
   {
      const int   c  = slotGlobal[l];
      HbCellInfo& ci = cells[l];
      fillRegions( ci, c );
      // all cells sharing at least one vertex, sorted by global id, including the cell itself
      std::set< int > nb;
      for ( int j = 0; j < 4; j++ )
         for ( int o : topo.vCells[topo.cellV[c][j]] )
            nb.insert( o );
      ci.copyBegin = (int) copies.size();
      for ( int o : nb )
      {
         HbCopy cp{};
         auto   it = globalToSlot.find( o );
         if ( it == globalToSlot.end() )
         {
            it = globalToSlot.emplace( o, (int) slotGlobal.size() ).first;
            slotGlobal.push_back( o );
         }
         cp.slot       = it->second;
         cp.sharedMask = 0;
         for ( int j = 0; j < 4; j++ )
         {
            cp.vmap[j] = -1;
            for ( int k = 0; k < 4; k++ )
               if ( topo.cellV[o][k] == topo.cellV[c][j] )
               {
                  cp.vmap[j] = (int8_t) k;
                  cp.sharedMask |= (uint8_t) ( 1u << j );
               }
         }
         copies.push_back( cp );
      }
      ci.copyEnd = (int) copies.size();
      // per interface region: the macro-cells that contain the primitive, with the coordinate map into each of them
      for ( int m = 0; m <= HB_NMASK; m++ )
         ci.rcBegin[m] = (int) regionCopies.size();
      for ( int m = 1; m < HB_NMASK; m++ )
      {
         ci.rcBegin[m] = (int) regionCopies.size();
         if ( ci.region[m].dim > 2 )
            continue;
         // support = local vertices with non-zero barycentric weight on the region; b0 <-> bit3, b1 <-> bit0, b2 <-> bit1, b3 <-> bit2
         static const int faceBitOfB[4] = { 8, 1, 2, 4 };
         int              supp          = 0;
         for ( int j = 0; j < 4; j++ )
            if ( !( m & faceBitOfB[j] ) )
               supp |= 1 << j;
         for ( int e = ci.copyBegin; e < ci.copyEnd; e++ )
         {
            const HbCopy& cp = copies[e];
            if ( supp & ~cp.sharedMask )
               continue;
            HbRegionCopy rc{};
            rc.slot = cp.slot;
            // b = ( N-x-y-z, x, y, z ) as rows over ( N, x, y, z )
            static const int bRow[4][4] = { { 1, -1, -1, -1 }, { 0, 1, 0, 0 }, { 0, 0, 1, 0 }, { 0, 0, 0, 1 } };
            int              bp[4][4]   = { { 0 } };
            bool             nz[4]      = { false, false, false, false };
            for ( int j = 0; j < 4; j++ )
               if ( ( supp >> j ) & 1 )
               {
                  const int k = cp.vmap[j];
                  nz[k]       = true;
                  for ( int t = 0; t < 4; t++ )
                     bp[k][t] += bRow[j][t];
               }
            for ( int k = 1; k < 4; k++ )
               for ( int t = 0; t < 4; t++ )
                  rc.c[k - 1][t] = (int8_t) bp[k][t];
            rc.maskp = (uint8_t) ( ( nz[1] ? 0 : 1 ) | ( nz[2] ? 0 : 2 ) | ( nz[3] ? 0 : 4 ) | ( nz[0] ? 0 : 8 ) );
            regionCopies.push_back( rc );
         }
      }
      ci.rcBegin[HB_NMASK] = (int) regionCopies.size();
   }
   // region tables of the mirrored (remote) cells: needed for multiplicities in the grid transfer
   cells.resize( slotGlobal.size(), HbCellInfo{} );
   for ( size_t sl = (size_t) nLocal; sl < slotGlobal.size(); sl++ )
      fillRegions( cells[sl], slotGlobal[sl] );

   // ghost exchange plan
   sendPlan.clear();
   recvPlan.clear();
   if ( nRanks > 1 )
   {
      for ( int peer = 0; peer < nRanks; peer++ )
      {
         if ( peer == rank )
            continue;
         for ( int c = 0; c < topo.nC; c++ ) // ascending global id
         {
            if ( topo.cellRank[c] == rank )
            {
               const uint32_t m = sharedRegionMasks( topo, c, peer );
               if ( m )
                  sendPlan.push_back( ExchangeEntry{ peer, globalToSlot.at( c ), m } );
            }
            else if ( topo.cellRank[c] == peer )
            {
               const uint32_t m = sharedRegionMasks( topo, c, rank );
               if ( m )
                  recvPlan.push_back( ExchangeEntry{ peer, globalToSlot.at( c ), m } );
            }
         }
      }
   }
}

} // namespace hb
