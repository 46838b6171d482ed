// Host-callable launchers of the kernels in kernels.cu.
#pragma once
#include "hb_compat.h"
#include "hb_types.h"

#define HB_REDUCE_MAX_BLOCKS 1184 // 148 SMs x 8

struct HbStencilArgs;

// dst = sum_k c[k]*src[k] + shift (+ dst if accumulate); also used for the product of functions
struct HbLinComb
{
   double*       dst;
   const double* src[4];
   double        c[4];
   double        shift;
   int32_t       n;
   int32_t       accumulate;
   int32_t       flag;
   HbBC          bc;
};

void hb_launch_stencil( int mode, const HbLevelView& lv, const HbStencilArgs& a, bool interior, bool boundary, cudaStream_t st );
void hb_launch_commit( const HbLevelView& lv, const HbStencilArgs& a, cudaStream_t st );
void hb_launch_lincomb( const HbLevelView& lv, const HbLinComb& lc, cudaStream_t st );
void hb_launch_multelem( const HbLevelView& lv, const HbLinComb& lc, cudaStream_t st );
// op 0: sum a*b, 1: sum a, 2: max |a| over owned points selected by flag; result in *out (device)
void hb_launch_reduce( const HbLevelView& lv, const HbBC& bc, int flag, int op, const double* a, const double* b, double* partial, double* out, cudaStream_t st );
void hb_launch_restrict( const HbLevelView& coarse, const HbLevelView& fine, const double* const* fsrc, double* cdst, const HbBC& bc, int flag, cudaStream_t st );
void hb_launch_prolongate( const HbLevelView& coarse, const HbLevelView& fine, const double* csrc, double* fdst, const HbBC& bc, int flag, int add, cudaStream_t st );
// ghost exchange: buf[i] = arena[off[i]] / arena[off[i]] = buf[i]
void hb_launch_pack( const double* arena, const long long* off, long long n, double* buf, cudaStream_t st );
void hb_launch_unpack( double* arena, const long long* off, long long n, const double* buf, cudaStream_t st );
// whole SOR sweep of a small level in one single-CTA launch; false if the level is too large (or in the emulation build)
bool hb_launch_sor_small( const HbLevelView& lv, const HbStencilArgs& a, int backwards, cudaStream_t st );
// interface phases kLo..kHi (0 vertices | 1,2 edge colours | 3,4,5 face colours; + commits) of an SOR sweep in one cooperative
// launch; false if unavailable
bool hb_launch_sor_interface_coop( const HbLevelView& lv, const HbStencilArgs& a, int backwards, int kLo, int kHi, cudaStream_t st );
extern int hb_use_desc; // 1: interface kernel reads per-block descriptors, 0: walks the topology tables (default)
