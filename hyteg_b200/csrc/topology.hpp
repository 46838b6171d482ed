// Macro-mesh topology: the host-side input of the matrix-free path.
//
// Plays the role of HyTeG's SetupPrimitiveStorage/PrimitiveStorage for the hot path only
// (reference src/hyteg/primitivestorage/SetupPrimitiveStorage.cpp, src/hyteg/primitives/Cell.hpp:34-203):
// macro vertices/edges/faces/cells, their neighbourhood counts, mesh boundary flags and the rank
// each macro-cell is assigned to.  Local numbering conventions follow HyTeG:
//   local edges  0:(v0,v1) 1:(v0,v2) 2:(v1,v2) 3:(v0,v3) 4:(v1,v3) 5:(v2,v3)
//   local faces  0:(v0,v1,v2) 1:(v0,v1,v3) 2:(v0,v2,v3) 3:(v1,v2,v3)
#pragma once
#include <array>
#include <cstdint>
#include <map>
#include <vector>

#include "hb_types.h"

namespace hb {

struct Topology
{
   int nV = 0, nE = 0, nF = 0, nC = 0;
   std::vector< double >               xyz;   // 3*nV
   std::vector< std::array< int, 4 > > cellV; // macro-cell -> global vertex ids (mesh order)
   std::vector< std::array< int, 2 > > edgeV; // sorted vertex ids
   std::vector< std::array< int, 3 > > faceV; // sorted vertex ids
   std::vector< std::array< int, 6 > > cellE;
   std::vector< std::array< int, 4 > > cellF;
   std::vector< int > vNumCells, eNumCells, fNumCells;
   std::vector< int > vMinCell, eMinCell, fMinCell;
   std::vector< int > vFlag, eFlag, fFlag, cFlag; // meshBoundaryFlag per primitive
   std::vector< uint8_t > vOnBnd, eOnBnd, fOnBnd, cOnBnd;
   std::vector< int > cellRank;
   std::vector< std::vector< int > > vCells; // cells around each macro vertex (sorted)

   void build( int nVertices, const double* coords, int nCells, const int* cells );
   // reference SetupPrimitiveStorage::setMeshBoundaryFlagsOnBoundary (SetupPrimitiveStorage.cpp:977-988)
   void setMeshBoundaryFlagsOnBoundary( int flagOnBoundary, int flagInner, bool highestDimAlwaysInner );
   // round-robin-by-block partition of macro-cells over ranks (contiguous chunks keep interfaces small)
   void partition( int nRanks );

   // mask -> primitive (dim, id) for a cell
   void regionPrimitive( int cell, int mask, int& dim, int& id ) const;
};

// mask <-> local primitive helpers
extern const int kFaceOfMask[16];   // mask with one bit  -> local face id, else -1
extern const int kEdgeOfMask[16];   // mask with two bits -> local edge id, else -1
extern const int kVertexOfMask[16]; // mask with three bits -> local vertex id, else -1
extern const int kLocalEdgeV[6][2];
extern const int kLocalFaceV[4][3];

// What one rank sees: owned macro-cells first, then mirrored (remote) neighbour cells.
struct LocalView
{
   int                rank = 0, nRanks = 1;
   std::vector< int > slotGlobal;          // slot -> global cell id; [0,nLocal) owned
   int                nLocal = 0;
   std::map< int, int > globalToSlot;
   std::vector< HbCellInfo > cells;        // [nLocal] (region tables + copy ranges)
   std::vector< HbCopy >     copies;
   std::vector< HbRegionCopy > regionCopies;

   // Ghost exchange plan (n_ranks > 1), level independent part.  A rank needs, of a remote macro-cell g, the values within
   // stencil distance 1 of every macro primitive (face / edge / vertex) g shares with one of the rank's own cells:
   // 'masks' has bit m set for every such region m of g.  Sender and receiver derive the same entries in the same order
   // (by peer rank, then by global cell id) from the topology alone, so no handshake is needed.  This is the role of
   // HyTeG's communicate< Cell, Face >() ... ghost-layer updates (reference communication/BufferedCommunication.cpp:181-420).
   struct ExchangeEntry
   {
      int      peer;  // remote rank
      int      slot;  // send: owned slot, recv: mirrored slot
      uint32_t masks; // regions of the cell whose layers travel
   };
   std::vector< ExchangeEntry > sendPlan, recvPlan;

   void build( const Topology& topo, int rank, int nRanks );
};

// bit m set <=> some macro-cell owned by 'rank' contains the macro primitive 'region m' of macro-cell g
uint32_t sharedRegionMasks( const Topology& topo, int g, int rank );

} // namespace hb
