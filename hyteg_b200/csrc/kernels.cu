// Hand-written sm_100a kernels of the P1 constant-stencil multigrid path.
//
// Replaces (on the device) the pystencils-generated host kernels of the reference:
//   apply_3D_macrocell_vertexdof_to_vertexdof_{replace,add}   (P1generatedKernels/...replace.cpp:33-79)
//   apply_3D_macroface_one_sided_*, vertexdof::macroedge/macrovertex::apply (P1Operator.hpp:192-320)
//   sor_3D_macrocell_P1 / sor_3D_macroface_P1                 (sor_3D_macrocell_P1.cpp:32-81)
//   assign_3D_macrocell_vertexdof_*_rhs_function(s), add_*    (p1functionspace/generatedKernels)
//   restrict_3D_macrocell_P1_pull_additive, prolongate_3D_macrocell_P1_push_additive
//   and all "communicate"/"communicateAdditively" ghost-layer copies between primitives.
// See DESIGN.md for the data layout and the roofline of each kernel.
#include "kernels.h"

#include "hb_device.h"
#include <stdexcept>
#include <string>

#ifdef HB_EMUL
dim3 threadIdx, blockIdx, blockDim, gridDim;
#else
#include <cooperative_groups.h>
#endif

#define HB_WPB 8 // warps (z-slices) per interior block
#define HB_BLOCK ( HB_WPB * 32 )

// =============================================================================================
//  Stencil kernel: interior tiles (register marching along y) + boundary points in one launch
// =============================================================================================

struct HbRow3
{
   double l, c, r;
};

template < bool NC >
__device__ __forceinline__ double hb_ld( const double* p )
{
   // non-coherent (read-only) path only when the kernel never writes the array it reads
   return NC ? __ldg( p ) : *p;
}

template < bool NC >
__device__ __forceinline__ HbRow3 hb_load_row( const double* S, int base, int last, int x )
{
   // row occupies S[base .. base+last]; x >= 1
   HbRow3 v;
   v.l = ( x - 1 <= last ) ? hb_ld< NC >( S + base + x - 1 ) : 0.0;
   v.c = ( x <= last ) ? hb_ld< NC >( S + base + x ) : 0.0;
   v.r = ( x + 1 <= last ) ? hb_ld< NC >( S + base + x + 1 ) : 0.0;
   return v;
}

template < int MODE >
__device__ __forceinline__ void hb_interior_tile( const HbLevelView& lv, const HbStencilArgs& a, int cell, const HbTile t, int warp, int lane )
{
   const int N = lv.N;
   const int z = t.z0 + warp;
   if ( z > N - 3 )
      return;
   const int x     = t.x0 + lane;
   const int y0    = t.y0;
   int       yLast = y0 + lv.tileRows - 1;  // last row of this tile
   if ( yLast > N - 2 - z )
      yLast = N - 2 - z;
   if ( yLast > N - 1 - x - z )
      yLast = N - 1 - x - z;               // this lane is outside the tet beyond that row
   if ( yLast < y0 )
      return;

   const HbRegion reg = lv.cells[cell].region[0];
   if ( !( ( a.selMask >> reg.bflag ) & 1 ) )
      return;

   const double* S = a.src[cell];
   constexpr bool NC = ( MODE != HB_SOR );
   const double* __restrict__ Wc = a.W + (size_t) cell * HB_NMASK * HB_NDIR;
   double w[HB_NDIR];
#pragma unroll
   for ( int d = 0; d < HB_NDIR; d++ )
      w[d] = __ldg( Wc + d );

   const int wm     = N + 1 - z;                       // width of slice z
   const int total  = hb_tet( N + 1 );
   const int sliceM = total - hb_tet( wm );
   const int sliceB = total - hb_tet( wm + 1 );
   const int sliceT = total - hb_tet( wm - 1 );
   // row base inside a slice of width ws: y*(ws+1) - y(y+1)/2 ; last valid x: ws-1-y
#define HB_ROWBASE( slice, ws, yy ) ( ( slice ) + ( yy ) * ( ( ws ) + 1 ) - ( ( yy ) * ( ( yy ) + 1 ) ) / 2 )

   HbRow3 P  = hb_load_row< NC >( S, HB_ROWBASE( sliceM, wm, y0 - 1 ), wm - 1 - ( y0 - 1 ), x );
   HbRow3 C  = hb_load_row< NC >( S, HB_ROWBASE( sliceM, wm, y0 ), wm - 1 - y0, x );
   HbRow3 B0 = hb_load_row< NC >( S, HB_ROWBASE( sliceB, wm + 1, y0 ), wm - y0, x );
   HbRow3 T0 = hb_load_row< NC >( S, HB_ROWBASE( sliceT, wm - 1, y0 - 1 ), wm - 2 - ( y0 - 1 ), x );

   const size_t cellOff = (size_t) cell * lv.cellSize;
   for ( int y = y0; y <= yLast; y++ )
   {
      const HbRow3 Nn = hb_load_row< NC >( S, HB_ROWBASE( sliceM, wm, y + 1 ), wm - 1 - ( y + 1 ), x );
      const HbRow3 B1 = hb_load_row< NC >( S, HB_ROWBASE( sliceB, wm + 1, y + 1 ), wm - ( y + 1 ), x );
      const HbRow3 T1 = hb_load_row< NC >( S, HB_ROWBASE( sliceT, wm - 1, y ), wm - 2 - y, x );

      double s = w[D_W] * C.l;
      s += w[D_E] * C.r;
      s += w[D_N] * Nn.c;
      s += w[D_NW] * Nn.l;
      s += w[D_S] * P.c;
      s += w[D_SE] * P.r;
      s += w[D_BC] * B0.c;
      s += w[D_BE] * B0.r;
      s += w[D_BN] * B1.c;
      s += w[D_BNW] * B1.l;
      s += w[D_TC] * T1.c;
      s += w[D_TW] * T1.l;
      s += w[D_TS] * T0.c;
      s += w[D_TSE] * T0.r;

      const size_t gi = cellOff + (size_t) ( HB_ROWBASE( sliceM, wm, y ) + x );
      if ( MODE != HB_SOR || ( ( x + 2 * y + 3 * z ) & 3 ) == a.colour )
         hb_epilogue< MODE >( a, a.dst + gi, gi, s, w[D_C], C.c );

      P  = C;
      C  = Nn;
      B0 = B1;
      T0 = T1;
   }
#undef HB_ROWBASE
}

// One block = one chunk (<= 256 points of ONE interface region) of one macro-cell.  Only the rank-local owner of the macro
// primitive (lowest macro-cell id on this rank) evaluates the stencil; the result goes to every copy of the point the rank holds
// (all copies of a DoF hold identical values, so the owner's centre / rhs / dst operands stand for all of them).
template < int MODE >
__device__ __forceinline__ void hb_boundary_block( const HbLevelView& lv, const HbStencilArgs& a, int blk )
{
   const int cell = blk / lv.nChunks;
   if ( cell >= lv.nCells )
      return;
   const HbBndChunk  ch  = lv.chunks[blk - cell * lv.nChunks];
   const HbCellInfo* ci  = lv.cells + cell;
   const HbRegion    reg = ci->region[ch.mask];
   if ( !( reg.owner & 2 ) )
      return;
   if ( !( ( a.selMask >> reg.bflag ) & 1 ) )
      return;
   if ( !( ( a.dimSel >> reg.dim ) & 1 ) )
      return;
   // dimSel bit 4: only primitives whose macro-cells all live on this rank; bit 5: only primitives shared with other ranks
   if ( ( ( a.dimSel & 16 ) && ( reg.owner & 4 ) ) || ( ( a.dimSel & 32 ) && !( reg.owner & 4 ) ) )
      return;
   if ( (int) threadIdx.x >= ch.count )
      return;
   const int        i = ch.first + threadIdx.x;
   const HbBndPoint p = lv.bnd[i];
   if ( a.bndColour >= 0 && hb_interface_colour( *ci, lv.N, p.x, p.y, p.z ) != a.bndColour )
      return;
   double sNb, diag;
   hb_boundary_stencil( lv, a.W, a.src, a.srcBase, ci, ch.mask, p.x, p.y, p.z, sNb, diag );
   const size_t local  = (size_t) hb_index( lv.N, p.x, p.y, p.z );
   const size_t gi     = (size_t) cell * lv.cellSize + local;
   const double centre = a.srcBase[gi];
   double       val;
   hb_epilogue< MODE >( a, &val, gi, sNb, diag, centre ); // CHEB_STEP also updates aux[gi]
   if ( a.toScratch )
   {
      a.scratch[(size_t) cell * lv.nBnd + i] = val; // fanned out by hb_commit_kernel
      return;
   }
   const double auxVal = ( MODE == HB_CHEB_STEP ) ? a.aux[gi] : 0.0;
   const int    e0 = ci->rcBegin[ch.mask], e1 = ci->rcBegin[ch.mask + 1];
   for ( int e = e0; e < e1; e++ )
   {
      const HbRegionCopy rc = lv.regionCopies[e];
      if ( rc.slot >= lv.nCells )
         continue; // mirrored macro-cell of another rank: its owner computes the same value
      const size_t gj = (size_t) rc.slot * lv.cellSize + (size_t) hb_copy_index( lv.N, rc, p.x, p.y, p.z );
      a.dst[gj]       = val;
      if ( MODE == HB_CHEB_STEP )
         a.aux[gj] = auxVal;
   }
}

// grid = nInteriorBlocks + nBoundaryBlocks.  Interior block b -> (cell, tile); its 8 warps take 8
// consecutive z-slices of the same (y,x) tile so that the slices shared between neighbouring warps
// are served by L1 instead of L2.
template < int MODE >
__global__ void __launch_bounds__( HB_BLOCK )
    hb_stencil_kernel( const HbLevelView lv, const HbStencilArgs a, const int nInteriorBlocks, const int doBoundary )
{
   const int b = blockIdx.x;
   if ( b < nInteriorBlocks )
   {
      const int    cell = b / lv.nTiles;
      const HbTile t    = lv.tiles[b - cell * lv.nTiles];
      hb_interior_tile< MODE >( lv, a, cell, t, threadIdx.x >> 5, threadIdx.x & 31 );
   }
   else if ( doBoundary )
      hb_boundary_block< MODE >( lv, a, b - nInteriorBlocks );
}

// Interface points only (large levels: the interior runs in its own streaming kernel).  Small register budget so that four
// CTAs are resident per SM: the kernel is a chain of dependent table look-ups and gathers, i.e. latency-bound.
template < int MODE >
__global__ void __launch_bounds__( HB_BLOCK, 4 ) hb_interface_kernel( const HbLevelView lv, const HbStencilArgs a )
{
   hb_boundary_block< MODE >( lv, a, blockIdx.x );
}

// scratch -> all local copies, for the interface phases of SOR
__global__ void hb_commit_kernel( const HbLevelView lv, const HbStencilArgs a )
{
   const long long item = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   const int       cell = (int) ( item / lv.nBnd );
   if ( cell >= lv.nCells )
      return;
   const int         i   = (int) ( item - (long long) cell * lv.nBnd );
   const HbBndPoint  p   = lv.bnd[i];
   const HbCellInfo* ci  = lv.cells + cell;
   const HbRegion    reg = ci->region[p.mask];
   if ( !( reg.owner & 2 ) )
      return;
   if ( !( ( a.selMask >> reg.bflag ) & 1 ) )
      return;
   if ( !( ( a.dimSel >> reg.dim ) & 1 ) )
      return;
   if ( a.bndColour >= 0 && hb_interface_colour( *ci, lv.N, p.x, p.y, p.z ) != a.bndColour )
      return;
   const double val = a.scratch[item];
   const int    e0 = ci->rcBegin[p.mask], e1 = ci->rcBegin[p.mask + 1];
   for ( int e = e0; e < e1; e++ )
   {
      const HbRegionCopy rc = lv.regionCopies[e];
      if ( rc.slot >= lv.nCells )
         continue;
      a.dst[(size_t) rc.slot * lv.cellSize + (size_t) hb_copy_index( lv.N, rc, p.x, p.y, p.z )] = val;
   }
}

#ifndef HB_EMUL
// =============================================================================================
//  Ghost exchange over NVLink peer memory
// =============================================================================================
// The ack word only says "the lines of message n have been read": the reads completed (their values were stored to the arena and a
// barrier passed) before the ack store is issued, and the waiting side orders nothing but its own later stores after the poll --
// so plain volatile (relaxed, L1-bypassing) accesses do; a release store here would stall on the thread's outstanding NVLink puts.
__device__ __forceinline__ unsigned long long hb_ld_word( const unsigned long long* p )
{
   unsigned long long v;
   asm volatile( "ld.volatile.global.u64 %0, [%1];" : "=l"( v ) : "l"( p ) : "memory" );
   return v;
}
__device__ __forceinline__ void hb_st_word( unsigned long long* p, unsigned long long v )
{
   asm volatile( "st.volatile.global.u64 [%0], %1;" ::"l"( p ), "l"( v ) : "memory" );
}
__device__ __forceinline__ bool hb_p2p_sends( const HbP2PArgs& p, int k ) { return p.sendStart[k + 1] > p.sendStart[k]; }
__device__ __forceinline__ bool hb_p2p_recvs( const HbP2PArgs& p, int k ) { return p.recvStart[k + 1] > p.recvStart[k]; }
#define HB_P2P_SPIN_LIMIT ( 1u << 27 ) // polls before a waiting kernel gives up (trap) instead of hanging the GPU: > 10 s

// Wire format: every double travels as one 16-byte line { lo, tag, hi, tag } (tag = low 32 bits of the pair's message number),
// written with one vector store; the receiver polls the line itself until both tags match.  Data and "it has landed" arrive
// together -- no system fence and no separate flag round trip on the critical path.  (8-byte single-copy atomicity of the two
// { half, tag } pairs is all this needs; the scheme is the one NCCL's LL protocol uses.)
__device__ __forceinline__ void hb_p2p_store_line( HbP2PLine* line, double v, unsigned tag )
{
   const unsigned long long b = (unsigned long long) __double_as_longlong( v );
   asm volatile( "st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"( line ), "r"( (unsigned) b ), "r"( tag ), "r"( (unsigned) ( b >> 32 ) ), "r"( tag ) : "memory" );
}
__device__ __forceinline__ double hb_p2p_load_line( const HbP2PLine* line, unsigned tag )
{
   unsigned lo, t0, hi, t1, spins = 0;
   do
   {
      asm volatile( "ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"( lo ), "=r"( t0 ), "=r"( hi ), "=r"( t1 ) : "l"( line ) : "memory" );
      if ( ++spins > HB_P2P_SPIN_LIMIT )
         __trap(); // the neighbour never sent this message (ranks out of step, or a rank died)
   } while ( t0 != tag || t1 != tag );
   return __longlong_as_double( (long long) ( ( (unsigned long long) hi << 32 ) | lo ) );
}

// message m of this launch may go into its window slot once message m-2 of the pair (same slot) has been unpacked over there
__device__ __forceinline__ void hb_p2p_wait_slot( const HbP2PArgs& p, unsigned m )
{
   const int k = threadIdx.x;
   if ( k < p.np && hb_p2p_sends( p, k ) && p.sendSeq[k] + m > 2 )
   {
      unsigned spins = 0;
      while ( hb_ld_word( p.localAck[k] ) + 2 < p.sendSeq[k] + m )
         if ( ++spins > HB_P2P_SPIN_LIMIT )
            __trap();
   }
   __syncthreads();
}
// A message carries the groups gLo..gHi of every peer's range (HbP2PArgs::sendGrp / recvGrp): element e of peer k's part sits at
// sendOff[ sendStart[k] + sendGrp[k][gLo] + e ] in the tables and on line e of the pair's window slot.  Pairs with an empty part still
// count the message (sequence numbers and acks advance for every pair that exchanges anything at all).
__device__ __forceinline__ void hb_p2p_put( const double* arena, const long long* __restrict__ sendOff, const HbP2PArgs& p, unsigned m, int gLo, int gHi,
                                            long long t0, long long stride )
{
   long long cum[HB_P2P_MAXPEERS + 1];
   cum[0] = 0;
   for ( int k = 0; k < p.np; k++ )
      cum[k + 1] = cum[k] + ( p.sendGrp[k][gHi + 1] - p.sendGrp[k][gLo] );
   const long long nS = cum[p.np];
   for ( long long i = t0; i < nS; i += stride )
   {
      int k = 0;
      while ( i >= cum[k + 1] )
         k++;
      const long long          e    = i - cum[k];
      const unsigned long long seq  = p.sendSeq[k] + m;
      HbP2PLine*               slot = p.remoteData[k] + ( ( seq & 1 ) ? p.segLines : 0 );
      hb_p2p_store_line( slot + e, arena[sendOff[p.sendStart[k] + p.sendGrp[k][gLo] + e]], (unsigned) seq );
   }
}
__device__ __forceinline__ void hb_p2p_get( double* arena, const long long* __restrict__ recvOff, const HbP2PArgs& p, unsigned m, int gLo, int gHi,
                                            long long t0, long long stride )
{
   long long cum[HB_P2P_MAXPEERS + 1];
   cum[0] = 0;
   for ( int k = 0; k < p.np; k++ )
      cum[k + 1] = cum[k] + ( p.recvGrp[k][gHi + 1] - p.recvGrp[k][gLo] );
   const long long nR = cum[p.np];
   for ( long long j = t0; j < nR; j += stride )
   {
      int k = 0;
      while ( j >= cum[k + 1] )
         k++;
      const long long          e    = j - cum[k];
      const unsigned long long seq  = p.recvSeq[k] + m;
      const HbP2PLine*         slot = p.localData[k] + ( ( seq & 1 ) ? p.segLines : 0 );
      arena[recvOff[p.recvStart[k] + p.recvGrp[k][gLo] + e]] = hb_p2p_load_line( slot + e, (unsigned) seq );
   }
}
// one CTA, after every get of the grid is done: the slot may be overwritten (by message m+2)
__device__ __forceinline__ void hb_p2p_ack( const HbP2PArgs& p, unsigned m )
{
   const int k = threadIdx.x;
   if ( k < p.np && hb_p2p_recvs( p, k ) )
      hb_st_word( p.remoteAck[k], p.recvSeq[k] + m );
}

// exchange number m of a launch from INSIDE a kernel whose CTAs can synchronise through bar() (block barrier of a single-CTA
// kernel, grid barrier of a cooperative launch): everything written to the arena before the call travels
template < class Bar >
__device__ __forceinline__ void hb_p2p_exchange_inline( double* arena, const HbP2PArgs& p, unsigned m, int gLo, int gHi, Bar bar )
{
   const long long t0 = (long long) blockIdx.x * blockDim.x + threadIdx.x, stride = (long long) gridDim.x * blockDim.x;
   bar();
   hb_p2p_wait_slot( p, m );
   hb_p2p_put( arena, p.sendOff, p, m, gLo, gHi, t0, stride );
   hb_p2p_get( arena, p.recvOff, p, m, gLo, gHi, t0, stride );
   bar(); // (no __threadfence() here: it would wait for this thread's NVLink puts to be acknowledged; the barrier orders the arena writes)
   if ( blockIdx.x == 0 )
      hb_p2p_ack( p, m );
}

// Ghost exchange in one kernel (replaces pack -> ncclSend/ncclRecv -> unpack; same data, same layout):
//   put : arena[sendOff[i]] -> lines in the neighbour's window (remote stores over NVLink)
//   get : poll the lines the neighbours write here, -> arena[recvOff[j]]; the last CTA acks
// Windows are double-buffered per pair.  No CTA waits on another CTA of its own grid, so the grid need not be co-resident; a
// neighbour's put never waits on this rank's current kernel.
__global__ void __launch_bounds__( 256 ) hb_p2p_exchange_kernel( double* arena, const HbP2PArgs p )
{
   __shared__ int  last;
   const long long stride = (long long) gridDim.x * blockDim.x;
   const long long t0     = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   hb_p2p_wait_slot( p, 0 );
   hb_p2p_put( arena, p.sendOff, p, 0, p.gLo, p.gHi, t0, stride );
   hb_p2p_get( arena, p.recvOff, p, 0, p.gLo, p.gHi, t0, stride );
   __syncthreads();
   if ( threadIdx.x == 0 )
      last = atomicAdd( p.counters, 1u ) == gridDim.x - 1;
   __syncthreads();
   if ( last )
   {
      hb_p2p_ack( p, 0 );
      if ( threadIdx.x == 0 )
         p.counters[0] = 0;
   }
}
#endif

int hb_p2p_grid_waves = 0; // tuning "p2p_grid_waves": CTAs of the stand-alone exchange kernel, in units of 148 (0: default)
int hb_p2p_overlapped = 0; // set by the caller while the exchange is issued beside the interior kernel
void hb_launch_p2p_exchange( double* arena, const HbP2PArgs& p, cudaStream_t st )
{
#ifndef HB_EMUL
   long long nS = 0, nR = 0;
   for ( int k = 0; k < p.np; k++ )
   {
      nS += p.sendGrp[k][p.gHi + 1] - p.sendGrp[k][p.gLo];
      nR += p.recvGrp[k][p.gHi + 1] - p.recvGrp[k][p.gLo];
   }
   const long long n    = nS > nR ? nS : nR;
   long long       grid = ( n + 1023 ) / 1024; // 4 elements per thread
   if ( grid < 1 )
      grid = 1;
   // every element is a chain of dependent accesses (offset -> value -> remote store; poll -> offset -> store): what moves the data
   // is the number of threads in flight, and no CTA waits for another one of its grid, so the grid may exceed one wave
   // (measured on 2 x B200, 10.7 MB per direction: alone 96 us on 148 CTAs, 72 us on 4 x 148 and more; beside the interior kernel of an
   // apply the step takes 847 us with 148 CTAs and 890 us with 8 x 148 -- there the exchange should leave the SMs to the interior kernel)
   const long long cap = 148ll * ( hb_p2p_grid_waves > 0 ? hb_p2p_grid_waves : ( hb_p2p_overlapped ? 1 : 8 ) );
   if ( grid > cap )
      grid = cap;
   hb_p2p_exchange_kernel<<< (unsigned) grid, 256, 0, st >>>( arena, p );
#endif
}

// Whole SOR sweep of a SMALL level in one launch: a single CTA walks the phases of hb_operator_smooth_sor (macro vertices,
// edges in 2 colours, faces in 3 colours, each staged through scratch and committed; then the 4 interior colours) with a
// block barrier in between.  The coarse levels of a V-cycle are launch-latency bound: 16 launches per sweep become one.
// Same arithmetic, in the same order, as the multi-launch path (hb_boundary_block / hb_interior_tile).
#ifndef HB_EMUL
__global__ void __launch_bounds__( 1024 ) hb_sor_small_kernel( const HbLevelView lv, HbStencilArgs a, const int backwards, const HbP2PArgs px )
{
   // more than one rank (px.np > 0): the ghost exchanges of the sweep happen in here as well, over the peer-memory windows --
   // after every interface phase but the last one, and (backward sweep) between the interior and the interface phases
   unsigned m        = 0;
   auto     exchange = [&]( int gLo, int gHi ) { // the point classes written since the last exchange (HB_NGROUP)
      if ( px.np > 0 )
         hb_p2p_exchange_inline( a.dst, px, m++, gLo, gHi, [] { __syncthreads(); } );
   };
   const int N = lv.N;
   // the phases walk precomputed work lists (HbLevelView::ifaceItems / phaseItems): no per-phase scan of all interface points
   auto interfacePhase = [&]( int dim, int colour ) {
      const int k  = dim == 0 ? 0 : ( dim == 1 ? 1 + colour : 3 + colour );
      const int i0 = lv.ifaceStart[k], i1 = lv.ifaceStart[k + 1];
      for ( int pass = 0; pass < 2; pass++ ) // 0: evaluate into scratch, 1: commit to all local copies
      {
         for ( int it = i0 + threadIdx.x; it < i1; it += blockDim.x )
         {
            const HbIfaceItem q = lv.ifaceItems[it];
            if ( !( ( a.selMask >> q.bflag ) & 1 ) )
               continue;
            if ( pass == 0 )
            {
               double sNb, diag;
               hb_boundary_stencil_range( lv, a.W, a.src, a.srcBase, q.rc0, q.rc1, q.x, q.y, q.z, sNb, diag );
               const size_t gi = (size_t) q.cell * lv.cellSize + (size_t) hb_index( N, q.x, q.y, q.z );
               double       val;
               hb_epilogue< HB_SOR >( a, &val, gi, sNb, diag, a.srcBase[gi] );
               a.scratch[it] = val;
            }
            else
            {
               const double val = a.scratch[it];
               for ( int e = q.rc0; e < q.rc1; e++ )
               {
                  const HbRegionCopy rc = lv.regionCopies[e];
                  if ( rc.slot < lv.nCells )
                     a.dst[(size_t) rc.slot * lv.cellSize + (size_t) hb_copy_index( N, rc, q.x, q.y, q.z )] = val;
               }
            }
         }
         __syncthreads();
      }
   };
   auto interiorPhase = [&]( int colour ) {
      for ( int it = lv.phaseStart[6 + colour] + threadIdx.x; it < lv.phaseStart[7 + colour]; it += blockDim.x )
      {
         const HbPhaseItem q = lv.phaseItems[it];
         if ( !( ( a.selMask >> q.bflag ) & 1 ) )
            continue;
         const int     cell = q.cell, x = q.x, y = q.y, z = q.z, local = hb_index( N, x, y, z );
         const size_t  g    = (size_t) cell * lv.cellSize + local;
         const double* w    = a.W + (size_t) cell * HB_NMASK * HB_NDIR;
         const double* S    = a.srcBase + (size_t) cell * lv.cellSize;
         // same summation order as hb_interior_tile
         double sum = w[D_W] * S[hb_index( N, x - 1, y, z )];
         sum += w[D_E] * S[hb_index( N, x + 1, y, z )];
         sum += w[D_N] * S[hb_index( N, x, y + 1, z )];
         sum += w[D_NW] * S[hb_index( N, x - 1, y + 1, z )];
         sum += w[D_S] * S[hb_index( N, x, y - 1, z )];
         sum += w[D_SE] * S[hb_index( N, x + 1, y - 1, z )];
         sum += w[D_BC] * S[hb_index( N, x, y, z - 1 )];
         sum += w[D_BE] * S[hb_index( N, x + 1, y, z - 1 )];
         sum += w[D_BN] * S[hb_index( N, x, y + 1, z - 1 )];
         sum += w[D_BNW] * S[hb_index( N, x - 1, y + 1, z - 1 )];
         sum += w[D_TC] * S[hb_index( N, x, y, z + 1 )];
         sum += w[D_TW] * S[hb_index( N, x - 1, y, z + 1 )];
         sum += w[D_TS] * S[hb_index( N, x, y - 1, z + 1 )];
         sum += w[D_TSE] * S[hb_index( N, x + 1, y - 1, z + 1 )];
         hb_epilogue< HB_SOR >( a, a.dst + g, g, sum, w[D_C], S[local] );
      }
      __syncthreads();
   };
   const int nColours[3] = { 1, 2, 3 };
   if ( px.syncFirst )
      exchange( px.gLo, px.gHi ); // catch-up: whatever was stale when the sweep started
   if ( !backwards )
   {
      for ( int d = 0; d <= 2; d++ )
         for ( int c = 0; c < nColours[d]; c++ )
         {
            interfacePhase( d, c );
            const int g = d == 0 ? 0 : ( d == 1 ? 1 + c : 3 + c );
            if ( !( d == 2 && c == 2 ) )
               exchange( g, g );
         }
      for ( int c = 0; c < 4; c++ )
         interiorPhase( c );
   }
   else
   {
      for ( int c = 3; c >= 0; c-- )
         interiorPhase( c );
      exchange( px.gLo, px.gHi ); // the interior class and whatever was stale when the sweep started
      for ( int d = 2; d >= 0; d-- )
         for ( int c = nColours[d] - 1; c >= 0; c-- )
         {
            interfacePhase( d, c );
            const int g = d == 1 ? 1 + c : 3 + c;
            if ( d != 0 )
               exchange( g, g );
         }
   }
}
#endif

// All interface phases of one SOR sweep (macro vertices, edges in 2 colours, faces in 3 colours; each evaluated into scratch,
// then committed) in ONE cooperative launch: persistent CTAs walk the (macro-cell, chunk) blocks of every phase and meet at a
// grid barrier between the passes, instead of 12 launches.  Same arithmetic as hb_interface_kernel< HB_SOR > + hb_commit_kernel.
#ifndef HB_EMUL
__device__ __forceinline__ void hb_commit_block( const HbLevelView& lv, const HbStencilArgs& a, int blk )
{
   const int cell = blk / lv.nChunks;
   if ( cell >= lv.nCells )
      return;
   const HbBndChunk  ch  = lv.chunks[blk - cell * lv.nChunks];
   const HbCellInfo* ci  = lv.cells + cell;
   const HbRegion    reg = ci->region[ch.mask];
   if ( !( reg.owner & 2 ) || !( ( a.selMask >> reg.bflag ) & 1 ) || !( ( a.dimSel >> reg.dim ) & 1 ) )
      return;
   if ( (int) threadIdx.x >= ch.count )
      return;
   const int        i = ch.first + threadIdx.x;
   const HbBndPoint p = lv.bnd[i];
   if ( a.bndColour >= 0 && hb_interface_colour( *ci, lv.N, p.x, p.y, p.z ) != a.bndColour )
      return;
   const double val = a.scratch[(size_t) cell * lv.nBnd + i];
   for ( int e = ci->rcBegin[ch.mask]; e < ci->rcBegin[ch.mask + 1]; e++ )
   {
      const HbRegionCopy rc = lv.regionCopies[e];
      if ( rc.slot < lv.nCells )
         a.dst[(size_t) rc.slot * lv.cellSize + (size_t) hb_copy_index( lv.N, rc, p.x, p.y, p.z )] = val;
   }
}

// One interior point of an SOR colour pass in the association order of the shuffle kernel (left column, centre column seeded
// with it, right column added last; shuffle_kernel.cu) -- bit-identical to it, so that a level may run its interior colours
// through either kernel.
__device__ __forceinline__ void hb_sor_interior_point( const HbLevelView& lv, const HbStencilArgs& a, const HbPhaseItem q )
{
   if ( !( ( a.selMask >> q.bflag ) & 1 ) )
      return;
   const int     N = lv.N, x = q.x, y = q.y, z = q.z;
   const double* w = a.W + (size_t) q.cell * HB_NMASK * HB_NDIR;
   const double* S = a.srcBase + (size_t) q.cell * lv.cellSize;
   auto          V = [&]( int dx, int dy, int dz ) { return S[hb_index( N, x + dx, y + dy, z + dz )]; };
   double l = w[D_W] * V( -1, 0, 0 );
   l        = fma( w[D_NW], V( -1, 1, 0 ), l );
   l        = fma( w[D_BNW], V( -1, 1, -1 ), l );
   l        = fma( w[D_TW], V( -1, 0, 1 ), l );
   const double centre = V( 0, 0, 0 );
   double       c      = fma( w[D_C], centre, l );
   c                   = fma( w[D_N], V( 0, 1, 0 ), c );
   c                   = fma( w[D_S], V( 0, -1, 0 ), c );
   c                   = fma( w[D_BC], V( 0, 0, -1 ), c );
   c                   = fma( w[D_BN], V( 0, 1, -1 ), c );
   c                   = fma( w[D_TC], V( 0, 0, 1 ), c );
   c                   = fma( w[D_TS], V( 0, -1, 1 ), c );
   double r = w[D_E] * V( 1, 0, 0 );
   r        = fma( w[D_SE], V( 1, -1, 0 ), r );
   r        = fma( w[D_BE], V( 1, 0, -1 ), r );
   r        = fma( w[D_TSE], V( 1, -1, 1 ), r );
   const size_t g = (size_t) q.cell * lv.cellSize + (size_t) hb_index( N, x, y, z );
   a.dst[g]       = hb_sor_value( a.omega, c + r, 1.0 / w[D_C], centre, a.rhs[g] );
}

// writes the new value of an interface point into every local macro-cell array that holds a copy of it
__device__ __forceinline__ void hb_iface_commit( const HbLevelView& lv, const HbStencilArgs& a, const HbIfaceItem& q, double val )
{
   for ( int e = q.rc0; e < q.rc1; e++ )
   {
      const HbRegionCopy rc = lv.regionCopies[e];
      if ( rc.slot < lv.nCells )
         a.dst[(size_t) rc.slot * lv.cellSize + (size_t) hb_copy_index( lv.N, rc, q.x, q.y, q.z )] = val;
   }
}

__global__ void __launch_bounds__( HB_BLOCK, 4 )
    hb_sor_interface_coop_kernel( const HbLevelView lv, HbStencilArgs a, const int backwards, const int kLo, const int kHi, const HbP2PArgs px,
                                  const int interior )
{
   cooperative_groups::grid_group grid = cooperative_groups::this_grid();
   unsigned                       m    = 0; // more than one rank (px.np > 0): ghost exchange after every phase but the sweep's last
   const int nBlk = lv.nCells * lv.nChunks;
   a.toScratch    = 1;
   // interior != 0 (levels with interior work lists, N <= 64): the four interior colours run in here as well -- the whole sweep
   // is one launch.  Forward: after the interface phases; backward: before them, followed by a ghost exchange.
   auto interiorColours = [&]() {
      const int t0 = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
      for ( int cc = 0; cc < 4; cc++ )
      {
         const int c = backwards ? 3 - cc : cc;
         for ( int it = lv.phaseStart[6 + c] + t0; it < lv.phaseStart[7 + c]; it += stride )
            hb_sor_interior_point( lv, a, lv.phaseItems[it] );
         if ( cc < 3 || backwards )
            grid.sync();
      }
   };
   if ( interior && backwards )
   {
      interiorColours();
      if ( px.np > 0 )
         hb_p2p_exchange_inline( a.dst, px, m++, px.gLo, px.gHi, [&] { grid.sync(); } ); // interior class + what was stale before the sweep
   }
   for ( int ph = 0; ph < 6; ph++ )
   {
      const int k   = backwards ? 5 - ph : ph;          // phases 0: vertices | 1,2: edges | 3,4,5: faces
      if ( k < kLo || k > kHi )
         continue;
      const int dim = k == 0 ? 0 : ( k <= 2 ? 1 : 2 );
      a.dimSel      = 1 << dim;
      a.bndColour   = k == 0 ? 0 : ( k <= 2 ? k - 1 : k - 3 );
      if ( lv.ifaceItems && ( k <= 2 || lv.ifaceFaces ) )
      {
         // list form: one load tells a thread its point, the macro-cells containing it and the primitive's boundary flag.
         // Only the points with a same-phase neighbour on another primitive (HbIfaceItem::stage, the first ifaceStaged[k] items
         // of the list) go through scratch; all others are written to their copies at once.
         const int i0 = lv.ifaceStart[k], i1 = lv.ifaceStart[k + 1], iS = i0 + lv.ifaceStaged[k];
         const int t0 = i0 + blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
         for ( int it = t0; it < i1; it += stride )
         {
            const HbIfaceItem q = lv.ifaceItems[it];
            if ( !( ( a.selMask >> q.bflag ) & 1 ) )
               continue;
            double sNb, diag;
            hb_boundary_stencil_range( lv, a.W, a.src, a.srcBase, q.rc0, q.rc1, q.x, q.y, q.z, sNb, diag );
            const size_t gi = (size_t) q.cell * lv.cellSize + (size_t) hb_index( lv.N, q.x, q.y, q.z );
            double       val;
            hb_epilogue< HB_SOR >( a, &val, gi, sNb, diag, a.srcBase[gi] );
            if ( q.stage )
               a.scratch[it] = val;
            else
               hb_iface_commit( lv, a, q, val );
         }
         if ( iS > i0 )
         {
            grid.sync();
            for ( int it = t0; it < iS; it += stride )
            {
               const HbIfaceItem q = lv.ifaceItems[it];
               if ( ( a.selMask >> q.bflag ) & 1 )
                  hb_iface_commit( lv, a, q, a.scratch[it] );
            }
         }
      }
      else
      {
         for ( int blk = blockIdx.x; blk < nBlk; blk += gridDim.x )
            hb_boundary_block< HB_SOR >( lv, a, blk );
         grid.sync();
         for ( int blk = blockIdx.x; blk < nBlk; blk += gridDim.x )
            hb_commit_block( lv, a, blk );
      }
      if ( px.np > 0 && k != ( backwards ? 0 : 5 ) )
         hb_p2p_exchange_inline( a.dst, px, m++, k, k, [&] { grid.sync(); } ); // phase k wrote point class k; starts and ends with a grid barrier
      else
         grid.sync();
   }
   if ( interior && !backwards )
      interiorColours();
}
#endif

#ifndef HB_EMUL
static int hb_coop_per_sm = -1, hb_coop_sms = 0;
#endif
bool hb_sor_coop_available()
{
#ifdef HB_EMUL
   return false;
#else
   if ( hb_coop_per_sm < 0 )
   {
      int dev = 0, coop = 0;
      cudaGetDevice( &dev );
      cudaDeviceGetAttribute( &hb_coop_sms, cudaDevAttrMultiProcessorCount, dev );
      cudaDeviceGetAttribute( &coop, cudaDevAttrCooperativeLaunch, dev );
      if ( !coop || cudaOccupancyMaxActiveBlocksPerMultiprocessor( &hb_coop_per_sm, hb_sor_interface_coop_kernel, HB_BLOCK, 0 ) != cudaSuccess )
         hb_coop_per_sm = 0;
   }
   return hb_coop_per_sm >= 1;
#endif
}

bool hb_launch_sor_interface_coop( const HbLevelView& lv, const HbStencilArgs& a, int backwards, int kLo, int kHi, const HbP2PArgs& px, int interior, cudaStream_t st )
{
#ifdef HB_EMUL
   return false;
#else
   if ( !hb_sor_coop_available() )
      return false;
   const int nSm = hb_coop_sms, perSm = hb_coop_per_sm;
   const int nBlk = lv.nCells * lv.nChunks;
   int       grid = nSm * ( hb_coop_ctas_per_sm > 0 && hb_coop_ctas_per_sm < perSm ? hb_coop_ctas_per_sm : perSm );
   if ( grid > nBlk )
      grid = nBlk;
   if ( grid < 1 )
      return true;
   HbLevelView   lvc = lv;
   HbStencilArgs ac  = a;
   int           bw  = backwards;
   HbP2PArgs     pxc = px;
   void*         args[] = { &lvc, &ac, &bw, &kLo, &kHi, &pxc, &interior };
   return cudaLaunchCooperativeKernel( (void*) hb_sor_interface_coop_kernel, dim3( grid ), dim3( HB_BLOCK ), args, 0, st ) == cudaSuccess;
#endif
}

bool hb_sor_small_fits( const HbLevelView& lv )
{
#ifdef HB_EMUL
   return false;
#else
   return (long long) lv.nCells * lv.cellSize <= 16384;
#endif
}

bool hb_launch_sor_small( const HbLevelView& lv, const HbStencilArgs& a, int backwards, const HbP2PArgs& px, cudaStream_t st )
{
#ifdef HB_EMUL
   return false;
#else
   if ( !hb_sor_small_fits( lv ) || !lv.phaseItems || !lv.ifaceItems || !lv.ifaceFaces )
      return false; // more than 16 strided passes per interior colour: the multi-launch path wins
   hb_sor_small_kernel<<< 1, 1024, 0, st >>>( lv, a, backwards, px );
   return true;
#endif
}

int hb_coop_ctas_per_sm = 0; // A/B knob "sor_coop_ctas": CTAs per SM of the cooperative SOR launch (0: as many as fit)

template < int MODE >
static void hb_launch_stencil_mode( const HbLevelView& lv, const HbStencilArgs& a, bool interior, bool boundary, cudaStream_t st )
{
   const long long nIntBlocks = interior ? (long long) lv.nCells * lv.nTiles : 0;
   const long long nBndBlocks = boundary ? (long long) lv.nCells * lv.nChunks : 0;
   const long long grid       = nIntBlocks + nBndBlocks;
   if ( grid == 0 )
      return;
   if ( nIntBlocks == 0 )
   {
      HB_LAUNCH( hb_interface_kernel< MODE >, (unsigned) grid, HB_BLOCK, st, lv, a );
      return;
   }
   HB_LAUNCH( hb_stencil_kernel< MODE >, (unsigned) grid, HB_BLOCK, st, lv, a, (int) nIntBlocks, boundary ? 1 : 0 );
}

void hb_launch_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, bool interior, bool boundary, cudaStream_t st )
{
   switch ( mode )
   {
   case HB_APPLY_REPLACE:
      hb_launch_stencil_mode< HB_APPLY_REPLACE >( lv, a, interior, boundary, st );
      break;
   case HB_APPLY_ADD:
      hb_launch_stencil_mode< HB_APPLY_ADD >( lv, a, interior, boundary, st );
      break;
   case HB_RESIDUAL:
      hb_launch_stencil_mode< HB_RESIDUAL >( lv, a, interior, boundary, st );
      break;
   case HB_JACOBI:
      hb_launch_stencil_mode< HB_JACOBI >( lv, a, interior, boundary, st );
      break;
   case HB_SOR:
      hb_launch_stencil_mode< HB_SOR >( lv, a, interior, boundary, st );
      break;
   case HB_CHEB_FIRST:
      hb_launch_stencil_mode< HB_CHEB_FIRST >( lv, a, interior, boundary, st );
      break;
   case HB_CHEB_STEP:
      hb_launch_stencil_mode< HB_CHEB_STEP >( lv, a, interior, boundary, st );
      break;
   case HB_DIAG:
      hb_launch_stencil_mode< HB_DIAG >( lv, a, interior, boundary, st );
      break;
   case HB_INVDIAG:
      hb_launch_stencil_mode< HB_INVDIAG >( lv, a, interior, boundary, st );
      break;
   }
}

void hb_launch_commit( const HbLevelView& lv, const HbStencilArgs& a, cudaStream_t st )
{
   const long long n = (long long) lv.nCells * lv.nBnd;
   if ( n == 0 )
      return;
   HB_LAUNCH( hb_commit_kernel, (unsigned) ( ( n + 255 ) / 256 ), 256, st, lv, a );
}

// =============================================================================================
//  BLAS-1 (flat over the local arena; boundary mask template decides DoFType / ownership)
// =============================================================================================

__device__ __forceinline__ bool hb_point_selected( const HbLevelView& lv, const HbBC& bc, int flag, size_t gi, bool ownerOnly )
{
   const int cell  = (int) ( gi / (size_t) lv.cellSize );
   const int local = (int) ( gi - (size_t) cell * lv.cellSize );
   const int m     = lv.mask[local];
   const HbRegion r = lv.cells[cell].region[m];
   if ( ownerOnly && !( r.owner & 1 ) )
      return false;
   return ( bc.type[r.bflag] & flag ) != 0;
}

// dst = sum_k c[k]*src[k]  (+ dst if accumulate)
__global__ void hb_lincomb_kernel( const HbLevelView lv, const HbLinComb lc )
{
   const size_t n      = (size_t) lv.nCells * lv.cellSize;
   const size_t stride = (size_t) gridDim.x * blockDim.x;
   for ( size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
   {
      if ( lc.flag != HB_DOF_ALL && !hb_point_selected( lv, lc.bc, lc.flag, i, false ) )
         continue;
      double v = lc.accumulate ? lc.dst[i] : 0.0;
      if ( lc.n > 0 )
         v += lc.c[0] * lc.src[0][i];
      if ( lc.n > 1 )
         v += lc.c[1] * lc.src[1][i];
      if ( lc.n > 2 )
         v += lc.c[2] * lc.src[2][i];
      if ( lc.n > 3 )
         v += lc.c[3] * lc.src[3][i];
      v += lc.shift;
      lc.dst[i] = v;
   }
}

// dst = prod_k src[k]
__global__ void hb_multelem_kernel( const HbLevelView lv, const HbLinComb lc )
{
   const size_t n      = (size_t) lv.nCells * lv.cellSize;
   const size_t stride = (size_t) gridDim.x * blockDim.x;
   for ( size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
   {
      if ( lc.flag != HB_DOF_ALL && !hb_point_selected( lv, lc.bc, lc.flag, i, false ) )
         continue;
      double v = lc.src[0][i];
      for ( int k = 1; k < lc.n; k++ )
         v *= lc.src[k][i];
      lc.dst[i] = v;
   }
}

// partial[b] = reduction over owned points: op 0: sum a*b, 1: sum a, 2: max |a|
__global__ void hb_reduce_kernel( const HbLevelView lv, const HbBC bc, int flag, int op, const double* a, const double* b, double* partial )
{
   const size_t n      = (size_t) lv.nCells * lv.cellSize;
   const size_t stride = (size_t) gridDim.x * blockDim.x;
   double       acc    = 0.0;
   for ( size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
   {
      if ( !hb_point_selected( lv, bc, flag, i, true ) )
         continue;
      if ( op == 0 )
         acc += a[i] * b[i];
      else if ( op == 1 )
         acc += a[i];
      else
         acc = fmax( acc, fabs( a[i] ) );
   }
#ifdef HB_EMUL
   if ( threadIdx.x == 0 )
      partial[blockIdx.x] = acc;
   else if ( op == 2 )
      partial[blockIdx.x] = fmax( partial[blockIdx.x], acc );
   else
      partial[blockIdx.x] += acc;
#else
   __shared__ double sm[32];
   for ( int o = 16; o > 0; o >>= 1 )
   {
      const double other = __shfl_down_sync( 0xffffffffu, acc, o );
      acc                = ( op == 2 ) ? fmax( acc, other ) : acc + other;
   }
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   if ( lane == 0 )
      sm[warp] = acc;
   __syncthreads();
   if ( warp == 0 )
   {
      const int nw = blockDim.x >> 5;
      acc          = lane < nw ? sm[lane] : 0.0;
      for ( int o = 16; o > 0; o >>= 1 )
      {
         const double other = __shfl_down_sync( 0xffffffffu, acc, o );
         acc                = ( op == 2 ) ? fmax( acc, other ) : acc + other;
      }
      if ( lane == 0 )
         partial[blockIdx.x] = acc;
   }
#endif
}

// out[slot] = reduction of partial[0..n) in fixed order (single block)
__global__ void hb_reduce_final_kernel( const double* partial, int n, int op, double* out )
{
#ifdef HB_EMUL
   double acc = 0.0;
   for ( int i = 0; i < n; i++ )
      acc = ( op == 2 ) ? fmax( acc, partial[i] ) : acc + partial[i];
   *out = acc;
#else
   __shared__ double sm[32];
   double            acc = 0.0;
   for ( int i = threadIdx.x; i < n; i += blockDim.x )
      acc = ( op == 2 ) ? fmax( acc, partial[i] ) : acc + partial[i];
   for ( int o = 16; o > 0; o >>= 1 )
   {
      const double other = __shfl_down_sync( 0xffffffffu, acc, o );
      acc                = ( op == 2 ) ? fmax( acc, other ) : acc + other;
   }
   const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
   if ( lane == 0 )
      sm[warp] = acc;
   __syncthreads();
   if ( warp == 0 )
   {
      const int nw = blockDim.x >> 5;
      acc          = lane < nw ? sm[lane] : 0.0;
      for ( int o = 16; o > 0; o >>= 1 )
      {
         const double other = __shfl_down_sync( 0xffffffffu, acc, o );
         acc                = ( op == 2 ) ? fmax( acc, other ) : acc + other;
      }
      if ( lane == 0 )
         *out = acc;
   }
#endif
}

static int hb_flat_grid( size_t n, int block, int maxBlocks )
{
   size_t g = ( n + block - 1 ) / block;
   if ( g > (size_t) maxBlocks )
      g = maxBlocks;
   if ( g < 1 )
      g = 1;
   return (int) g;
}

void hb_launch_lincomb( const HbLevelView& lv, const HbLinComb& lc, cudaStream_t st )
{
   const size_t n = (size_t) lv.nCells * lv.cellSize;
   if ( n == 0 )
      return;
   HB_LAUNCH( hb_lincomb_kernel, hb_flat_grid( n, 256, 148 * 16 ), 256, st, lv, lc );
}

void hb_launch_multelem( const HbLevelView& lv, const HbLinComb& lc, cudaStream_t st )
{
   const size_t n = (size_t) lv.nCells * lv.cellSize;
   if ( n == 0 )
      return;
   HB_LAUNCH( hb_multelem_kernel, hb_flat_grid( n, 256, 148 * 16 ), 256, st, lv, lc );
}

void hb_launch_reduce( const HbLevelView& lv, const HbBC& bc, int flag, int op, const double* a, const double* b, double* partial, double* out, cudaStream_t st )
{
   const size_t n = (size_t) lv.nCells * lv.cellSize;
#ifdef HB_EMUL
   const int block = 1;
#else
   const int block = 256;
#endif
   const int grid = hb_flat_grid( n, 256, HB_REDUCE_MAX_BLOCKS );
   HB_LAUNCH( hb_reduce_kernel, grid, block, st, lv, bc, flag, op, a, b, partial );
   HB_LAUNCH( hb_reduce_final_kernel, 1, block, st, partial, grid, op, out );
}

// ---------------------------------------------------------------------------------------------
//  Allocation-free reduction (hb_p1function_dot_no_alloc): the same two kernels as hb_launch_reduce, with the per-CTA partial sums in a
//  file-scope device array instead of a cudaMalloc'ed buffer.  HB_REDUCE_MAX_BLOCKS doubles (9.25 KB), resident on every GPU that loads the
//  library until the process exits.  In host code the array's NAME is only a host-side stand-in; the device address comes from
//  cudaGetSymbolAddress, resolved once per device with that device current and cached.  All calls share the array: nothing in it has to
//  start at zero (every slot the final pass reads is written by the first pass of the same call), but two calls that overlap on the GPU
//  (different streams) race on it.
// ---------------------------------------------------------------------------------------------
#ifdef HB_EMUL
static double hb_no_alloc_scratch[HB_REDUCE_MAX_BLOCKS];
#else
static __device__ __align__( 256 ) double hb_no_alloc_scratch[HB_REDUCE_MAX_BLOCKS];
#endif

static double* hb_no_alloc_scratch_address()
{
#ifdef HB_EMUL
   return hb_no_alloc_scratch;
#else
   constexpr int  MAXDEV         = 64;
   static double* cached[MAXDEV] = {};
   int            dev            = -1;
   if ( cudaGetDevice( &dev ) != cudaSuccess || dev < 0 || dev >= MAXDEV )
      throw std::runtime_error( "hyteg_b200: dot_no_alloc: no current CUDA device" );
   if ( !cached[dev] )
   {
      double*           ptr = nullptr;
      const cudaError_t e   = cudaGetSymbolAddress( (void**) &ptr, hb_no_alloc_scratch );
      if ( e != cudaSuccess || !ptr )
         throw std::runtime_error( std::string( "hyteg_b200: dot_no_alloc: cudaGetSymbolAddress failed: " ) + cudaGetErrorString( e ) );
      cached[dev] = ptr;
   }
   return cached[dev];
#endif
}

void hb_launch_reduce_no_alloc( const HbLevelView& lv, const HbBC& bc, int flag, int op, const double* a, const double* b, double* out, cudaStream_t st )
{
   const size_t n = (size_t) lv.nCells * lv.cellSize;
   if ( n > (size_t) HB_NO_ALLOC_MAX_ENTRIES )
      throw std::runtime_error( "hyteg_b200: dot_no_alloc: " + std::to_string( n ) + " entries on this level, the allocation-free path takes at most " +
                                std::to_string( (long long) HB_NO_ALLOC_MAX_ENTRIES ) + " (HB_NO_ALLOC_MAX_ENTRIES)" );
#ifdef HB_EMUL
   const int block = 1;
#else
   const int block = 256;
#endif
   const int grid = hb_flat_grid( n, 256, HB_REDUCE_MAX_BLOCKS ); // <= the size of the scratch array by construction
   double*   partial = hb_no_alloc_scratch_address();
   HB_LAUNCH( hb_reduce_kernel, grid, block, st, lv, bc, flag, op, a, b, partial );
   HB_LAUNCH( hb_reduce_final_kernel, 1, block, st, partial, grid, op, out );
}

// =============================================================================================
//  Ghost exchange: pack the layers a peer needs into a contiguous send buffer / unpack a received buffer into the
//  mirrored macro-cells (reference: VertexDoFPackInfo::packDataForMacroFace / unpack..., communication/BufferedCommunication.cpp)
// =============================================================================================

__global__ void hb_pack_kernel( const double* arena, const long long* off, long long n, double* buf )
{
   const long long stride = (long long) gridDim.x * blockDim.x;
   for ( long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
      buf[i] = arena[off[i]];
}

__global__ void hb_unpack_kernel( double* arena, const long long* off, long long n, const double* buf )
{
   const long long stride = (long long) gridDim.x * blockDim.x;
   for ( long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
      arena[off[i]] = buf[i];
}

void hb_launch_pack( const double* arena, const long long* off, long long n, double* buf, cudaStream_t st )
{
   if ( n > 0 )
      HB_LAUNCH( hb_pack_kernel, hb_flat_grid( (size_t) n, 256, 148 * 8 ), 256, st, arena, off, n, buf );
}

void hb_launch_unpack( double* arena, const long long* off, long long n, const double* buf, cudaStream_t st )
{
   if ( n > 0 )
      HB_LAUNCH( hb_unpack_kernel, hb_flat_grid( (size_t) n, 256, 148 * 8 ), 256, st, arena, off, n, buf );
}

// =============================================================================================
//  Grid transfer (gather form)
// =============================================================================================

// Full-weighting restriction fine -> coarse = transpose of linear interpolation.
// reference P1toP1LinearRestriction::restrict3D (P1toP1LinearRestriction.cpp:169-346) and
// restrict_3D_macrocell_P1_pull_additive: leaves on a macro face/edge/vertex are shared between
// macro-cells and therefore divided by the number of cells sharing that primitive; the additive
// communication of the reference becomes the sum over the copies of the coarse point.
__device__ __forceinline__ double hb_restrict_partial( const HbLevelView& fine, const double* f, int slot, int xc, int yc, int zc )
{
   // coarse point (xc,yc,zc) in macro-cell 'slot'; fine level N = 2*coarse N
   const int        Nf = fine.N;
   const int        x = 2 * xc, y = 2 * yc, z = 2 * zc;
   const int        m = hb_mask( Nf, x, y, z );
   const int        vm = hb_valid[m];
   const HbRegion*  reg = fine.cells[slot].region;
   double           acc = f[hb_index( Nf, x, y, z )] / (double) reg[m].mult;
#pragma unroll
   for ( int d = 1; d < HB_NDIR; d++ )
      if ( ( vm >> d ) & 1 )
      {
         const int xx = x + hb_dir[d][0], yy = y + hb_dir[d][1], zz = z + hb_dir[d][2];
         const int mm = hb_mask( Nf, xx, yy, zz );
         acc += 0.5 * f[hb_index( Nf, xx, yy, zz )] / (double) reg[mm].mult;
      }
   return acc;
}

// interface coarse point (cell, x, y, z): sum over the macro-cells containing it (additive communication of the reference)
__device__ __forceinline__ double hb_restrict_interface( const HbLevelView& coarse, const HbLevelView& fine, const double* const* fsrc, const HbCellInfo* ci, int x, int y, int z )
{
   const int N    = coarse.N;
   const int b[4] = { N - x - y - z, x, y, z };
   int       supp = 0;
   for ( int j = 0; j < 4; j++ )
      supp |= ( b[j] != 0 ) << j;
   double acc = 0.0;
   for ( int e = ci->copyBegin; e < ci->copyEnd; e++ )
   {
      const HbCopy cp = coarse.copies[e];
      if ( supp & ~cp.sharedMask )
         continue;
      int bp[4] = { 0, 0, 0, 0 };
      for ( int j = 0; j < 4; j++ )
         if ( ( supp >> j ) & 1 )
            bp[cp.vmap[j]] = b[j];
      acc += hb_restrict_partial( fine, fsrc[cp.slot], cp.slot, bp[1], bp[2], bp[3] );
   }
   return acc;
}

// pass 1: interior coarse points (flat over the arena; interface points are skipped -- or, with doInterface, handled in place:
// emulation build and tiny levels).  pass 2 (hb_restrict_interface_kernel): the compact interface point list, so that the long
// interface path does not drag every warp of pass 1 along (each row of the tetrahedral layout starts and ends on an interface).
__global__ void hb_restrict_kernel( const HbLevelView coarse, const HbLevelView fine, const double* const* fsrc, double* cdst, const HbBC bc, int flag, int doInterface )
{
   const size_t n      = (size_t) coarse.nCells * coarse.cellSize;
   const size_t stride = (size_t) gridDim.x * blockDim.x;
   for ( size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
   {
      const int         cell  = (int) ( i / (size_t) coarse.cellSize );
      const int         local = (int) ( i - (size_t) cell * coarse.cellSize );
      const HbBndPoint  q     = coarse.xyz[local];
      const int         x = q.x, y = q.y, z = q.z;
      const int         m  = q.mask;
      const HbCellInfo* ci = coarse.cells + cell;
      if ( m != 0 && !doInterface )
         continue;
      if ( !( bc.type[ci->region[m].bflag] & flag ) )
         continue;
      double acc = 0.0;
      if ( m == 0 )
      {
         // interior coarse point: the fine point 2(x,y,z) and its 14 neighbours are interior points of the same macro-cell
         // (multiplicity 1), so hb_restrict_partial reduces to f[c] + 0.5 * sum f[c + d] -- same order, no divisions
         const double* __restrict__ f = fsrc[cell];
         const int Nf = fine.N, xf = 2 * x, yf = 2 * y, zf = 2 * z;
         const int ws   = Nf + 1 - zf;
         const int i0   = hb_index( Nf, xf, yf, zf );
         const int szUp = ( ws * ( ws + 1 ) ) / 2, szDn = ( ( ws + 1 ) * ( ws + 2 ) ) / 2;
         const int dN = ws - yf, dS = -( ws - yf + 1 ), dT = szUp - yf, dTS = szUp - ws, dB = yf - szDn, dBN = ws + 1 - szDn;
         acc = f[i0];
         acc += 0.5 * f[i0 - 1];        // W
         acc += 0.5 * f[i0 + 1];        // E
         acc += 0.5 * f[i0 + dN];       // N
         acc += 0.5 * f[i0 + dS];       // S
         acc += 0.5 * f[i0 + dN - 1];   // NW
         acc += 0.5 * f[i0 + dS + 1];   // SE
         acc += 0.5 * f[i0 + dT];       // TC
         acc += 0.5 * f[i0 + dT - 1];   // TW
         acc += 0.5 * f[i0 + dTS];      // TS
         acc += 0.5 * f[i0 + dTS + 1];  // TSE
         acc += 0.5 * f[i0 + dB];       // BC
         acc += 0.5 * f[i0 + dBN];      // BN
         acc += 0.5 * f[i0 + dB + 1];   // BE
         acc += 0.5 * f[i0 + dBN - 1];  // BNW
      }
      else
         acc = hb_restrict_interface( coarse, fine, fsrc, ci, x, y, z );
      cdst[i] = acc;
   }
}

__global__ void hb_restrict_interface_kernel( const HbLevelView coarse, const HbLevelView fine, const double* const* fsrc, double* cdst, const HbBC bc, int flag )
{
   const long long n      = (long long) coarse.nCells * coarse.nBnd;
   const long long stride = (long long) gridDim.x * blockDim.x;
   for ( long long item = (long long) blockIdx.x * blockDim.x + threadIdx.x; item < n; item += stride )
   {
      const int         cell = (int) ( item / coarse.nBnd );
      const HbBndPoint  p    = coarse.bnd[item - (long long) cell * coarse.nBnd];
      const HbCellInfo* ci   = coarse.cells + cell;
      if ( !( bc.type[ci->region[p.mask].bflag] & flag ) )
         continue;
      cdst[(size_t) cell * coarse.cellSize + hb_index( coarse.N, p.x, p.y, p.z )] = hb_restrict_interface( coarse, fine, fsrc, ci, p.x, p.y, p.z );
   }
}

// Linear prolongation coarse -> fine, gather form: every fine point is either a coarse point or the
// midpoint of exactly one coarse micro-edge, whose direction is given by the parity of (x,y,z).
// reference P1toP1LinearProlongation::prolongate3DAdditively (P1toP1LinearProlongation.cpp:194-410).
__global__ void hb_prolongate_kernel( const HbLevelView coarse, const HbLevelView fine, const double* csrc, double* fdst, const HbBC bc, int flag, int add )
{
   const size_t n      = (size_t) fine.nCells * fine.cellSize;
   const size_t stride = (size_t) gridDim.x * blockDim.x;
   const int    Nf = fine.N, Nc = coarse.N;
   for ( size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride )
   {
      const int cell  = (int) ( i / (size_t) fine.cellSize );
      const int local = (int) ( i - (size_t) cell * fine.cellSize );
      const HbBndPoint q = fine.xyz[local];
      const int        x = q.x, y = q.y, z = q.z;
      const int m = hb_mask( Nf, x, y, z );
      if ( !( bc.type[fine.cells[cell].region[m].bflag] & flag ) )
         continue;
      const double* c  = csrc + (size_t) cell * coarse.cellSize;
      const int     px = x & 1, py = y & 1, pz = z & 1;
      double        v;
      if ( !( px | py | pz ) )
         v = c[hb_index( Nc, x >> 1, y >> 1, z >> 1 )];
      else
      {
         // edge direction with that parity: (1,0,0) (0,1,0) (0,0,1) (1,-1,0) (1,0,-1) (0,1,-1) (1,-1,1)
         int dx = px, dy = py, dz = pz;
         if ( px && py )
            dy = -1;
         else if ( px && pz )
            dz = -1;
         else if ( py && pz )
            dz = -1;
         if ( px && py && pz )
         {
            dy = -1;
            dz = 1;
         }
         const int ax = ( x - dx ) >> 1, ay = ( y - dy ) >> 1, az = ( z - dz ) >> 1;
         const int bx = ( x + dx ) >> 1, by = ( y + dy ) >> 1, bz = ( z + dz ) >> 1;
         v            = 0.5 * c[hb_index( Nc, ax, ay, az )] + 0.5 * c[hb_index( Nc, bx, by, bz )];
      }
      fdst[i] = add ? fdst[i] + v : v;
   }
}

void hb_launch_restrict( const HbLevelView& coarse, const HbLevelView& fine, const double* const* fsrc, double* cdst, const HbBC& bc, int flag, cudaStream_t st )
{
   const size_t n = (size_t) coarse.nCells * coarse.cellSize;
   if ( n == 0 )
      return;
#ifdef HB_EMUL
   const bool twoPass = false;
#else
   const bool twoPass = coarse.N >= 16;
#endif
   HB_LAUNCH( hb_restrict_kernel, hb_flat_grid( n, 256, 148 * 16 ), 256, st, coarse, fine, fsrc, cdst, bc, flag, twoPass ? 0 : 1 );
   if ( twoPass )
      HB_LAUNCH( hb_restrict_interface_kernel, hb_flat_grid( (size_t) coarse.nCells * coarse.nBnd, 128, 148 * 16 ), 128, st, coarse, fine, fsrc, cdst, bc, flag );
}

void hb_launch_prolongate( const HbLevelView& coarse, const HbLevelView& fine, const double* csrc, double* fdst, const HbBC& bc, int flag, int add, cudaStream_t st )
{
   const size_t n = (size_t) fine.nCells * fine.cellSize;
   if ( n == 0 )
      return;
   HB_LAUNCH( hb_prolongate_kernel, hb_flat_grid( n, 256, 148 * 16 ), 256, st, coarse, fine, csrc, fdst, bc, flag, add );
}
