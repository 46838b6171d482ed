// P2 (vertexdof + edgedof) function storage and the fused on-the-fly elementwise operators: POD argument blocks and launchers.
//
// Layout contract (follows HyTeG): a P2 function on level L is a P1 function (vertex DoFs, hb_function) plus, per macro-cell, seven
// edge-DoF arrays -- orientations X, Y, Z, XY, XZ, YZ as tetrahedral arrays of width N = 2^L and XYZ of width N - 1, one behind the other
// (reference src/hyteg/edgedofspace/EdgeDoFIndexing.hpp:922-985, EdgeDoFOrientation.hpp:29-39).  Like the vertex arrays they cover the
// CLOSED macro-tet: an edge DoF on a macro face / edge is stored in every adjacent macro-cell and all copies always hold the same value.
#pragma once
#include "hb_compat.h"
#include "hb_types.h"

#define HB_P2_MAXQ 11 // quadrature points of the largest rule (Xiao-Gimbutas 4)

enum HbP2Form
{
   HB_P2_DIVKGRAD = 0, // int k grad(u) . grad(v), 11-point degree-4 rule (forms/form_hyteg_generated/p2/p2_div_k_grad_blending_q4.cpp)
   HB_P2_EPSILON  = 1, // int 2 mu eps(phi_j e_d) : eps(phi_i e_c), 4-point degree-2 rule (p2_epsilonvar_blending_q2.cpp)
   HB_P2_DIV      = 2, // P2 -> P1:  - int psi_i d_c phi_j           (p2_to_p1/p2_to_p1_div_blending_q2.cpp, same 4-point rule)
   HB_P2_DIVT     = 3, // P1 -> P2:  - int d_c phi_i psi_j           (p1_to_p2/p1_to_p2_divt_blending_q2.cpp)
   HB_P2_EPSILON_VECTOR = 4, // all nine epsilon blocks at once on a P2 vector function: y^c_i = int mu ( grad u + grad u^T )_c . grad phi_i
   HB_P2_STOKES   = 5  // the whole Taylor-Hood operator in one pass: A u + B^T p ; B u (epsilon A-block, div / divT as above)
};
__host__ __device__ __forceinline__ bool hb_p2_form_q2( int form )
{
   return form != HB_P2_DIVKGRAD;
}

__host__ __device__ __forceinline__ int hb_p2_tet( int w )
{
   return ( w * ( w + 1 ) * ( w + 2 ) ) / 6;
}
// entries of one macro-cell's edge-DoF block
__host__ __device__ __forceinline__ int hb_edge_cell_size( int N )
{
   return 6 * hb_p2_tet( N ) + hb_p2_tet( N - 1 );
}
// index inside a tetrahedral array of width w (entries x + y + z <= w - 1), reference indexing/MacroCellIndexing.hpp:44-52
__host__ __device__ __forceinline__ int hb_tet_index( int w, int x, int y, int z )
{
   const int wz = w - z;
   return hb_p2_tet( w ) - hb_p2_tet( wz ) + y * ( wz + 1 ) - ( y * ( y + 1 ) ) / 2 + x;
}
// edgedof::macrocell::index( level, x, y, z, orientation ), orientation 0..6 = X Y Z XY XZ YZ XYZ
__host__ __device__ __forceinline__ int hb_edge_index( int N, int o, int x, int y, int z )
{
   return o * hb_p2_tet( N ) + hb_tet_index( o < 6 ? N : N - 1, x, y, z );
}

// everything the element kernel needs about one macro-cell: 4 vertices and the IcosahedralShellMap in derived form
// F(x) = x r(x) / |x|,  r(x) = radRay + dr ( (x - ray) . n ) / volT      (reference geometry/IcosahedralShellMap.hpp:109-239)
struct HbP2CellGeom
{
   double v[4][3];
   double ray[3], n[3]; // n = (thr - ray) x (for - ray)
   double radRay, drOverVolT;
   int32_t bflag; // meshBoundaryFlag of the macro-cell
   int32_t pad;
};

struct HbP2View
{
   int32_t           N, nCells;
   int32_t           vCellSize, eCellSize; // entries per macro-cell: vertex array, edge block
   int32_t           nAnchors;             // tet(N): logical micro-cell indices per type
   int32_t           blending;             // 0 affine, 1 IcosahedralShellMap
   const HbBndPoint* anchors;              // [nAnchors] (x,y,z) of the anchor, mask unused
   const uint8_t*    vmask;                // [vCellSize] boundary mask of a vertex entry (HbLevelView::mask)
   const uint8_t*    emask;                // [eCellSize] boundary mask of an edge entry = AND of its end points' masks
   const HbCellInfo* cells;                // region tables (bflag / owner per mask)
   const HbP2CellGeom* geom;               // [nCells]
};

struct HbP2ElemArgs
{
   const double* srcV[3]; // components 1, 2 only for HB_P2_EPSILON_VECTOR; srcE null for HB_P2_DIVT (P1 source)
   const double* srcE[3];
   double*       accV[3]; // zeroed accumulators, same layout as the function; accE null for HB_P2_DIV (P1 destination)
   double*       accE[3];
   const double* srcP;    // HB_P2_STOKES: pressure (P1) source and accumulator
   double*       accP;
   const double* kV;   // coefficient as a P2 function (or null)
   const double* kE;
   const double* kq;   // coefficient sampled at the quadrature points [cell][type][anchor][q] (or null; null and no kV: k = 1)
   int32_t       form, compC, compD;
   int32_t       diagonal; // 1: accumulate the operator's diagonal instead of A u (computeDiagonalOperatorValues)
};

struct HbP2FinalizeArgs
{
   double*       dstV; // either part may be null: it is left alone then
   double*       dstE;
   const double* accV;
   const double* accE;
   double        beta; // dst = beta * dst + acc on selected DoFs
   uint32_t      selMask;
   int32_t       pad;
   const long long* grpStart; // interface groups: copies of one DoF, entry = arena offset (bit 62: edge arena)
   const long long* grpEntry;
   const uint8_t*   grpFlag;  // meshBoundaryFlag of the group's primitive
   long long        nGroups;
};

// dst = sum_k c[k] src[k] (+ dst) on the edge DoFs selected by flag; also constants (n = 0, shift); op 1: dst = prod_k src[k]; op 2: dst = 1 / dst
struct HbP2LinComb
{
   int32_t       op, pad0;
   double*       dst;
   const double* src[4];
   double        c[4];
   double        shift;
   int32_t       n, accumulate;
   uint32_t      selMask;
   int32_t       pad;
};

// P2toP2QuadraticProlongation / Restriction between level l (coarse view) and l + 1 (fine view)
struct HbP2TransferArgs
{
   const double* fineV;   // restriction: source; prolongation: destination (const cast away there)
   const double* fineE;
   double*       coarseV; // restriction: zeroed accumulators of the coarse level; prolongation: source
   double*       coarseE;
   uint32_t      selMask; // prolongation: fine DoFs of selected primitives are written
   int32_t       add;     // prolongateAndAdd
   const HbBndPoint* fineXyz; // [fine.vCellSize] logical coordinates of the fine level's vertex entries (HbLevelView::xyz)
};
void hb_p2_launch_prolongate( const HbP2View& coarse, const HbP2View& fine, const HbP2TransferArgs& a, cudaStream_t st );
// scatters every OWNED fine DoF into the coarse accumulators (cell-local); hb_p2_launch_finalize on the coarse level completes it
void hb_p2_launch_restrict( const HbP2View& coarse, const HbP2View& fine, const HbP2TransferArgs& a, cudaStream_t st );

void hb_p2_launch_elements( const HbP2View& pv, const HbP2ElemArgs& a, cudaStream_t st );
void hb_p2_launch_finalize( const HbP2View& pv, const HbP2FinalizeArgs& a, cudaStream_t st );
void hb_p2_launch_edge_lincomb( const HbP2View& pv, const HbP2LinComb& lc, cudaStream_t st );
// partial[blocks] <- sum over OWNED selected edge entries of a*b; returns the number of blocks used
int hb_p2_launch_edge_dot( const HbP2View& pv, uint32_t selMask, const double* a, const double* b, double* partial, int maxBlocks, cudaStream_t st );
// physical positions of the quadrature points [cell][type][anchor][q][3] (so that a host callback can be sampled there)
void hb_p2_launch_quad_points( const HbP2View& pv, int form, double* out, cudaStream_t st );
// F(x) and DF(x) of the cell's map at n points given in computational coordinates: out[n][12]
void hb_p2_launch_map_eval( const HbP2View& pv, int cell, const double* x, int n, double* out, cudaStream_t st );
int  hb_p2_num_quad_points( int form );
