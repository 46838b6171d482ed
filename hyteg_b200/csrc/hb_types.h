// POD types shared between the host runtime and the sm_100a kernels.
//
// Layout contract (follows HyTeG, reference src/hyteg/indexing/MacroCellIndexing.hpp:40-52 and
// src/hyteg/Levelinfo.hpp:36-131): a P1 function on refinement level L stores, per macro-cell, a
// tetrahedral array of (N+1)(N+2)(N+3)/6 doubles with N = 2^L, x fastest, then y, then z.  The
// array covers the CLOSED macro-tet, i.e. the points on the macro faces/edges/vertices are stored
// in every adjacent macro-cell (HyTeG keeps them as ghost copies there as well).
#pragma once
#include <stdint.h>

#define HB_NDIR 15
#define HB_NMASK 16

// DoFType bit values, reference src/hyteg/types/types.hpp:35-44
#define HB_DOF_NONE 0
#define HB_DOF_INNER 1
#define HB_DOF_DIRICHLET 2
#define HB_DOF_NEUMANN 4
#define HB_DOF_FREESLIP 8
#define HB_DOF_ALL 15

// Boundary mask of a micro-vertex (x,y,z) inside a macro-cell with N subdivisions:
//   bit0: x == 0   (local face 2 = {v0,v2,v3})
//   bit1: y == 0   (local face 1 = {v0,v1,v3})
//   bit2: z == 0   (local face 0 = {v0,v1,v2})
//   bit3: x+y+z == N (local face 3 = {v1,v2,v3})
// one bit -> macro-face, two bits -> macro-edge, three bits -> macro-vertex.

struct HbRegion
{
   uint8_t bflag;   // meshBoundaryFlag of the macro-primitive this region lies on (0..15)
   uint8_t owner;   // 1 if this macro-cell is the owning copy (lowest global cell id)
   uint8_t mult;    // number of macro-cells sharing the primitive
   uint8_t dim;     // 0 vertex, 1 edge, 2 face, 3 cell interior, 255 invalid mask
};

// One macro-cell contributing to the stencil of an interface region (macro face / edge / vertex) of another macro-cell:
// a point (x,y,z) of the region has the coordinates p' = c * (N,x,y,z) in macro-cell 'slot', where it lies on region maskp.
struct HbRegionCopy
{
   int32_t slot;    // index into cell pointer table / W table
   uint8_t maskp;   // boundary mask of the image point in 'slot' (constant over the region)
   int8_t  c[3][4]; // rows x', y', z'; columns N, x, y, z
   uint8_t pad[3];
};

// A run of <= 256 interface points of one region in the (mask-sorted) interface point list of a level
struct HbBndChunk
{
   int32_t first, count, mask, pad;
};

struct HbCellInfo
{
   HbRegion region[HB_NMASK];
   int32_t  rcBegin[HB_NMASK + 1]; // region copies of region m: [rcBegin[m], rcBegin[m+1]) (ascending global cell id)
   int32_t  copyBegin; // range in the copy list (sorted by global cell id, contains the cell itself)
   int32_t  copyEnd;
   int32_t  globalId;
   int32_t  pad;
   int32_t  gv[4]; // global macro-vertex ids of the local vertices (canonical interface colouring)
};

// One entry of a macro-cell's copy list: a macro-cell (slot index into the cell pointer table)
// sharing at least one macro-vertex, together with the local vertex renumbering.
struct HbCopy
{
   int32_t slot;       // index into cell pointer table / W table
   uint8_t sharedMask; // bit j set: local vertex j of the reference cell is also a vertex of 'slot'
   int8_t  vmap[4];    // local vertex j of reference cell -> local vertex of 'slot' (or -1)
   uint8_t pad[3];
};

struct HbTile
{
   int16_t z0, y0, x0, pad; // first slice / row / column of an interior tile
};

struct HbBndPoint
{
   uint16_t x, y, z, mask;
};

// Per-level geometry/topology tables living in device memory.  Passed to kernels by value.
// interior work lists of the single-CTA SOR sweep (small levels): phases 6..9 = interior colours (i = local array index);
// the interface phases 0..5 use the HbIfaceItem lists below
struct HbPhaseItem
{
   int32_t  cell;
   uint16_t x, y, z;
   uint8_t  bflag; // meshBoundaryFlag of the macro-cell
   uint8_t  pad;
};
#define HB_NPHASE 10
// work lists of the cooperative SOR interface launch: everything a thread needs to start on its point in ONE load (the table walk
// chunk -> region -> point -> colour -> copy range is five dependent L2 round trips per phase, and a grid barrier empties L1).
// Phases 0 | 1,2 always; the face phases 3,4,5 only on levels whose six phases run cooperatively (N <= 128).
struct HbIfaceItem
{
   int32_t  cell;
   int32_t  rc0, rc1; // range in regionCopies: the macro-cells containing the point
   uint16_t x, y, z;
   uint8_t  bflag;
   uint8_t  stage; // 1: a neighbour of the point lies on ANOTHER primitive of the same class and is updated in the same phase, so
                   // the new value is staged in scratch and committed after a barrier (block Jacobi between primitives of a class,
                   // P1Operator.hpp:399-411); 0: no other point of the phase reads this one -- it is written in place at once.
                   // Staged items come first in every phase's list (HbLevelView::ifaceStaged).
};

struct HbLevelView
{
   int32_t level;
   int32_t N;         // 2^level
   int32_t cellSize;  // (N+1)(N+2)(N+3)/6
   int32_t nCells;    // macro-cells owned by this rank (work items)
   int32_t nSlots;    // owned + mirrored remote cells (pointer-table length)
   int32_t nBnd;      // boundary points per macro-cell
   int32_t nTiles;    // interior tiles per macro-cell
   int32_t tileRows;  // rows marched per tile
   const HbBndPoint* bnd;   // [nBnd]     (identical for every macro-cell), sorted by mask
   const uint8_t*    mask;  // [cellSize] boundary mask per local array index
   const HbBndPoint* xyz;   // [cellSize] logical coordinates + mask per local array index
   const HbTile*     tiles; // [nTiles]
   const HbCellInfo* cells; // [nSlots] (copy ranges only valid for the first nCells)
   const HbCopy*     copies;
   const HbRegionCopy* regionCopies;
   const HbBndChunk*   chunks; // [nChunks] interface points grouped by region (identical for every macro-cell)
   int32_t             nChunks;
   int32_t             pad0;
   const void*         reserved0;
   const HbIfaceItem*  ifaceItems;
   int32_t             ifaceStart[7]; // phase k: [ifaceStart[k], ifaceStart[k+1]); ifaceFaces = 0: phases 3..5 are not listed
   int32_t             ifaceFaces;
   int32_t             ifaceStaged[6]; // phase k: the items [ifaceStart[k], ifaceStart[k] + ifaceStaged[k]) are staged
   const HbPhaseItem*  phaseItems; // null on levels too large for the single-CTA sweep
   int32_t             phaseStart[HB_NPHASE + 1];
   int32_t             pad1;
};

// BoundaryCondition of a function: meshBoundaryFlag -> DoFType
// (reference src/hyteg/boundary/BoundaryCondition.hpp, create0123BC()).
struct HbBC
{
   uint8_t type[16];
};
