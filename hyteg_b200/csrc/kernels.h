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
// the same without any device allocation: partial sums in a file-scope __device__ array (kernels.cu); levels of more than
// HB_NO_ALLOC_MAX_ENTRIES arena entries are rejected (std::runtime_error)
#define HB_NO_ALLOC_MAX_ENTRIES ( 1ll << 28 )
void hb_launch_reduce_no_alloc( const HbLevelView& lv, const HbBC& bc, int flag, int op, const double* a, const double* b, double* out, cudaStream_t st );
void hb_launch_restrict( const HbLevelView& coarse, const HbLevelView& fine, const double* const* fsrc, double* cdst, const HbBC& bc, int flag, cudaStream_t st );
void hb_launch_prolongate( const HbLevelView& coarse, const HbLevelView& fine, const double* csrc, double* fdst, const HbBC& bc, int flag, int add, cudaStream_t st );
// ghost exchange: buf[i] = arena[off[i]] / arena[off[i]] = buf[i]
void hb_launch_pack( const double* arena, const long long* off, long long n, double* buf, cudaStream_t st );
void hb_launch_unpack( double* arena, const long long* off, long long n, const double* buf, cudaStream_t st );
// ghost exchange over peer memory (NVLink): one kernel packs the layers straight into the neighbours' windows as 16-byte
// { lo, tag, hi, tag } lines and unpacks the lines the neighbours write here as they land.  All pointers are device addresses;
// "remote" ones are windows of other GPUs mapped through CUDA IPC (runtime_comm.inc).
#define HB_P2P_MAXPEERS 8
struct alignas( 16 ) HbP2PLine
{
   uint32_t lo, tag0, hi, tag1;
};
// exchange groups (classes of travelling points): 0 macro vertices | 1, 2 macro edges by interface colour | 3..5 macro faces by
// interface colour | 6 macro-cell interiors.  Group k < 6 is exactly what SOR interface phase k writes.
#define HB_NGROUP 7
struct HbP2PArgs
{
   int32_t             np;
   int32_t             syncFirst; // fused sweeps: exchange once before the first phase (the mirrored layers were stale)
   long long           sendStart[HB_P2P_MAXPEERS + 1], recvStart[HB_P2P_MAXPEERS + 1];
   const long long*    sendOff;                     // arena offsets of the layers that travel / that arrive (Exchange tables)
   const long long*    recvOff;
   long long           segLines;                    // distance between the two slots of a pair, in 16-byte lines
   HbP2PLine*              remoteData[HB_P2P_MAXPEERS]; // neighbour's window for this rank's messages (slot 0; message n uses slot n & 1)
   unsigned long long* remoteAck[HB_P2P_MAXPEERS];  // neighbour's "message n to this rank has been unpacked"
   const HbP2PLine*              localData[HB_P2P_MAXPEERS];
   const unsigned long long* localAck[HB_P2P_MAXPEERS];
   unsigned long long  sendSeq[HB_P2P_MAXPEERS], recvSeq[HB_P2P_MAXPEERS]; // 1-based numbers of the launch's first message
   unsigned int*       counters;                                           // [2], zero between launches
   int32_t             sendGrp[HB_P2P_MAXPEERS][HB_NGROUP + 1], recvGrp[HB_P2P_MAXPEERS][HB_NGROUP + 1]; // group boundaries inside a peer's range
   int32_t             gLo, gHi; // stand-alone exchange kernel: the groups gLo..gHi travel
};
void hb_launch_p2p_exchange( double* arena, const HbP2PArgs& p, cudaStream_t st );
extern int hb_p2p_grid_waves, hb_p2p_overlapped;
// whole SOR sweep of a small level in one single-CTA launch; false if the level is too large (or in the emulation build)
// (px.np > 0: more than one rank, the sweep's ghost exchanges run inside the kernel over the peer-memory windows)
bool hb_sor_small_fits( const HbLevelView& lv );
bool hb_launch_sor_small( const HbLevelView& lv, const HbStencilArgs& a, int backwards, const HbP2PArgs& px, cudaStream_t st );
// interface phases kLo..kHi (0 vertices | 1,2 edge colours | 3,4,5 face colours; + commits) of an SOR sweep in one cooperative
// launch; false if unavailable
bool hb_sor_coop_available();
// interior != 0 (needs HbLevelView::phaseItems): the four interior colours run in the same launch -- the whole sweep
bool hb_launch_sor_interface_coop( const HbLevelView& lv, const HbStencilArgs& a, int backwards, int kLo, int kHi, const HbP2PArgs& px, int interior, cudaStream_t st );
extern int hb_coop_ctas_per_sm; // cooperative SOR launch: CTAs per SM (0: as many as fit)
