// Host runtime objects behind the C-ABI handles.
#pragma once
#include <array>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "hb_compat.h"
#include "hb_device.h"
#include "kernels.h"
#include "stream_kernel.h"
#include "shuffle_kernel.h"
#include "topology.hpp"

namespace hb {

void         cudaCheck( cudaError_t e, const char* what );
cudaStream_t currentStream();
int          tuning( const std::string& key, int dflt );
bool         p2pUsable(); // peer-memory windows connected and not switched off (tuning "p2p")
bool         deviceAllReduce( double* dev, int n, int op ); // over the NCCL communicator, on the library's stream; false: no communicator
// temporarily redirects the library's stream (ghost exchange on the side stream)
struct StreamScope
{
   cudaStream_t saved;
   explicit StreamScope( cudaStream_t s );
   ~StreamScope();
};

template < typename T >
struct DeviceBuffer
{
   T*     ptr = nullptr;
   size_t n   = 0;
   DeviceBuffer() {}
   DeviceBuffer( const DeviceBuffer& )            = delete;
   DeviceBuffer& operator=( const DeviceBuffer& ) = delete;
   ~DeviceBuffer() { release(); }
   void release()
   {
      if ( ptr )
         cudaFree( ptr );
      ptr = nullptr;
      n   = 0;
   }
   void alloc( size_t count )
   {
      release();
      n = count;
      cudaCheck( cudaMalloc( (void**) &ptr, ( count ? count : 1 ) * sizeof( T ) ), "cudaMalloc" );
   }
   void upload( const std::vector< T >& h )
   {
      alloc( h.size() );
      if ( !h.empty() )
         cudaCheck( cudaMemcpy( ptr, h.data(), h.size() * sizeof( T ), cudaMemcpyHostToDevice ), "cudaMemcpy H2D" );
   }
};

struct LevelTables
{
   HbLevelView                view{};
   DeviceBuffer< HbBndPoint > bnd, xyz;
   DeviceBuffer< uint8_t >    mask;
   DeviceBuffer< HbTile >     tiles;
   DeviceBuffer< HbBndChunk > chunks;
   std::vector< HbBndChunk >  hChunks;
   std::vector< HbBndPoint >  hBnd;
   DeviceBuffer< HbIfaceItem > ifaceItems; // work lists of the cooperative SOR interface launch
   DeviceBuffer< HbPhaseItem > phaseItems; // work lists of the single-CTA SOR sweep, rebuilt with the topology tables
   DeviceBuffer< HbItem >     itemsShuffle; // work items of the centre-column + shuffle interior kernel
   DeviceBuffer< double >     scratch; // nCells*nBnd, interface staging (SOR)
   std::vector< HbBndPoint >  hXyz;    // host copy of the coordinate template

   // ghost exchange of this level (n_ranks > 1, built on first use): element offsets into a function arena
   // ( slot * cellSize + local index ), grouped by peer in the order of LocalView::sendPlan / recvPlan
   struct Exchange
   {
      bool                         built = false;
      std::vector< int >           peers;                // ranks this rank exchanges with (ascending)
      std::vector< long long >     sendStart, recvStart; // [nPeers+1] ranges in sendOff / recvOff
      // Inside a peer's range the entries are ordered by the CLASS of the travelling point -- group 0: macro vertices, 1-2: macro
      // edges (interface colour 0, 1), 3-5: macro faces (colour 0..2), 6: macro-cell interiors -- so that an exchange can carry only
      // the classes written since the last one (an SOR phase writes exactly one).  sendGrp[k][g]: first entry of group g relative
      // to sendStart[k]; [HB_NGROUP]: the peer's count.
      std::vector< std::array< int32_t, HB_NGROUP + 1 > > sendGrp, recvGrp;
      DeviceBuffer< long long >    sendOff, recvOff;
      DeviceBuffer< double >       sendBuf, recvBuf;
   } ex;
};

} // namespace hb

struct hb_storage
{
   hb::Topology                             topo;
   hb::LocalView                            view;
   hb::DeviceBuffer< HbCellInfo >           dCells;
   hb::DeviceBuffer< HbCopy >               dCopies;
   hb::DeviceBuffer< HbRegionCopy >         dRegionCopies;
   std::map< int, std::unique_ptr< hb::LevelTables > > levels;
   hb::DeviceBuffer< double >               reducePartial; // HB_REDUCE_MAX_BLOCKS
   hb::DeviceBuffer< double >               reduceOut;     // 8 slots
   double*                                  hReduce = nullptr; // pinned
   bool                                     tablesDirty = true;

   hb::LevelTables& level( int L );
   hb::LevelTables::Exchange& exchange( int L );
   void             buildPhaseItems( hb::LevelTables& lt );
   void             uploadTopologyTables();
   ~hb_storage();
};

struct hb_function
{
   hb_storage* st;
   int         minLevel, maxLevel;
   HbBC        bc;
   std::vector< hb::DeviceBuffer< double > >  data;    // per level: local arena
   std::vector< hb::DeviceBuffer< double* > > cellPtr; // per level: [nSlots] device pointers
   std::vector< unsigned char > mirrorsStale;          // per level: bit g set = the mirrored cells do not hold the owners' current
                                                       // values of point class g (exchange groups, LevelTables::Exchange)
   void touched( int L ) // call after every write to the level's arena
   {
      touchedGroups( L, ( 1u << HB_NGROUP ) - 1u );
   }
   void touchedGroups( int L, unsigned groups ) // ... after a write to points of these classes only (SOR phases)
   {
      check( L );
      mirrorsStale[L - minLevel] |= (unsigned char) groups;
   }
   bool stale( int L ) const
   {
      check( L );
      return st_nranks() > 1 && mirrorsStale[L - minLevel] != 0;
   }
   int  st_nranks() const;
   // arguments for nMessages consecutive ghost exchanges of level L over the peer-memory windows (advances the pair counters);
   // false -- and out.np = 0 -- on one rank or when the windows are not in use
   bool p2pPrepare( int L, int nMessages, HbP2PArgs& out );
   void syncMirrors( int L );                                // ghost exchange if stale (no-op on one rank)
   double*        arena( int L ) const;
   double* const* ptrs( int L ) const;
   void           check( int L ) const;
};

struct hb_operator
{
   hb_storage* st;
   int         minLevel, maxLevel;
   std::vector< hb::DeviceBuffer< double > > W;  // per level [nSlots][16][15]
   std::vector< std::vector< double > >      hW; // host copy
   const double* weights( int L ) const;
   // Interior weights in constant memory (shuffle kernel).  If (levels x local cells) fit the table, all levels are resident and the
   // entry of (level L, local cell c) is constBase(L) + c; otherwise ONE level is resident at a time (entry c) and a launch on another
   // level re-uploads it, stream-ordered from a pinned staging buffer (no host synchronisation).  More local cells than table
   // entries: the shuffle kernel is not used (a one-time message says so).
   bool    constUsable() const;
   bool    constAllLevels() const;
   int     constBase( int L ) const;
   int     constChunks() const;
   void    ensureConstWeights( int L, int chunk = 0 );
   double* stage = nullptr; // pinned: [levels][nLocal][16] weights, then [levels][nLocal] flags
   ~hb_operator();
};
