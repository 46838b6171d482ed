// Launcher of the centre-column + shuffle interior stencil kernel (shuffle_kernel.cu).
#pragma once
#include <vector>

#include "stream_kernel.h" // HbItem, HbStreamArgs, HB_ARENA_PAD

bool hb_shuffle_supported( int mode, const HbLevelView& lv );
// work items of one macro-cell for N subdivisions: 16 slices x 30 columns x <= ys rows (ys <= 0: default);
// order 0: sorted by decreasing work, 1: natural (z-major) order
void hb_shuffle_build_items( int N, std::vector< HbItem >& items, int ys = 0, int order = 0 );
// interior points only; the interface points are handled by hb_launch_stencil( ..., interior=false, boundary=true )
void hb_launch_shuffle_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, const HbStreamArgs& sa, cudaStream_t st, int depth = 0 );
// uploads n weight sets ( 15 weights + 1/w[0] each, 16 doubles ) and the macro-cells' meshBoundaryFlags to constant-memory entries [first, first+n)
void hb_shuffle_upload_weights( int first, int n, const double* w16, const uint8_t* bflag, cudaStream_t st );
