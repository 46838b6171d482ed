// Launcher of the TMA-streamed interior stencil kernel (stream_kernel.cu).
#pragma once
#include <vector>

#include "hb_device.h"
#include "kernels.h"

#define HB_CONST_W_ENTRIES 448 // (level, macro-cell) weight sets held in constant memory (56 KB)
#define HB_ARENA_PAD 128 // doubles appended to every function arena so that staged row segments may overrun

// one work item of the streamed kernel, identical for every macro-cell of a level
struct HbItem
{
   int16_t z0, x0; // first output slice / column
   int16_t ya;     // first output row
   int16_t nOut;   // output rows
};

struct HbStreamArgs
{
   const double* srcArena;   // arena of the function the stencil reads
   const double* extraArena; // arena of the fused epilogue operand (rhs / dst / x), or null
   const HbItem* items;
   int32_t       nItems;
   uint32_t      selMask; // bit f set: macro primitives with meshBoundaryFlag f are selected by (bc, flag)
   int32_t       constBase;                        // shuffle kernel: first entry of this level's weights in constant memory
   int32_t       sync;      // shuffle kernel: pace the warps of a CTA with a named barrier per unrolled block
   int32_t       cellMajor; // shuffle kernel: block id = cell * nItems + item (else item * nCells + cell)
   int32_t*      counter; // [2] work counter + finished-CTA counter, both zero between launches
   int32_t       cellFirst, cellCount; // shuffle kernel: the launch covers macro-cells [cellFirst, cellFirst + cellCount) (cellCount 0: all)
};

// true if 'mode' on this level runs through the streamed kernel (N >= 32 and a streamed mode)
bool hb_stream_supported( int mode, const HbLevelView& lv );
// items of one macro-cell for N subdivisions
void hb_stream_build_items( int N, std::vector< HbItem >& items );
// interior points only; the interface points are still handled by hb_launch_stencil( ..., interior=false, boundary=true )
void hb_launch_stream_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st );
