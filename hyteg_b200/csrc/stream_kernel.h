// Work items and launch arguments of the interior stencil kernel (shuffle_kernel.cu).
#pragma once
#include <vector>

#include "hb_device.h"
#include "kernels.h"

#define HB_CONST_W_ENTRIES 448 // (level, macro-cell) weight sets held in constant memory (56 KB)
#define HB_ARENA_PAD 128 // doubles appended to every function arena: lanes right of the diagonal read past the last row of the last macro-cell

// one work item of the interior kernel, identical for every macro-cell of a level
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
   uint32_t      selMask;   // bit f set: macro primitives with meshBoundaryFlag f are selected by (bc, flag)
   int32_t       constBase; // first entry of this level's weights in constant memory
   int32_t       cellMajor; // block id = cell * nItems + item (else item * nCells + cell)
   int32_t       cellFirst, cellCount; // the launch covers the macro-cells [cellFirst, cellFirst + cellCount), whose weights sit in the
                                       // constant-memory entries constBase .. (more local macro-cells than entries: several launches)
};
