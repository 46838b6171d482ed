// C-ABI implementation, part 1: runtime, PrimitiveStorage tables, P1Function storage and BLAS-1.
#include "runtime.hpp"

#include <algorithm>
#include <cstring>

#include "../../include/hyteg_b200.h"
#include "p2_kernels.h"
#include "stencil_assembly.hpp"

namespace hb {

static thread_local std::string g_lastError;
static cudaStream_t             g_stream = 0;
static std::map< std::string, int > g_tuning;
hb_allreduce_fn                 g_allreduce    = nullptr;
void*                           g_allreduceCtx = nullptr;

void setError( const std::string& s )
{
   g_lastError = s;
}

void cudaCheck( cudaError_t e, const char* what )
{
   if ( e != cudaSuccess )
      throw std::runtime_error( std::string( "hyteg_b200: CUDA error in " ) + what + ": " + cudaGetErrorString( e ) );
}

cudaStream_t currentStream()
{
   return g_stream;
}

StreamScope::StreamScope( cudaStream_t s )
: saved( g_stream )
{
   g_stream = s;
}
StreamScope::~StreamScope()
{
   g_stream = saved;
}

int tuning( const std::string& key, int dflt )
{
   auto it = g_tuning.find( key );
   return it == g_tuning.end() ? dflt : it->second;
}

} // namespace hb

using namespace hb;

static const hb_operator* g_constOwner = nullptr; // operator whose interior weights + cell flags sit in constant memory

#define HB_TRY try {
#define HB_CATCH( ret )                   \
   }                                      \
   catch ( const std::exception& e )      \
   {                                      \
      hb::setError( e.what() );           \
      return ret;                         \
   }

static void require( bool cond, const char* msg )
{
   if ( !cond )
      throw std::runtime_error( std::string( "hyteg_b200: " ) + msg );
}

// ------------------------------------------------------------------------------------------------
// storage
// ------------------------------------------------------------------------------------------------

hb_storage::~hb_storage()
{
   if ( hReduce )
      cudaFreeHost( hReduce );
}

void hb_storage::uploadTopologyTables()
{
   view.build( topo, view.rank, view.nRanks );
   dCells.upload( view.cells );
   dCopies.upload( view.copies );
   dRegionCopies.upload( view.regionCopies );
   for ( auto& kv : levels )
   {
      buildPhaseItems( *kv.second );
      kv.second->view.cells  = dCells.ptr;
      kv.second->view.copies = dCopies.ptr;
      kv.second->view.regionCopies = dRegionCopies.ptr;
      kv.second->view.nSlots = (int) view.slotGlobal.size();
   }
   tablesDirty = false;
   g_constOwner = nullptr; // macro-cell flags may have changed (operators created afterwards stage the new flags)
}

void hb_storage::buildPhaseItems( hb::LevelTables& lt )
{
   HbLevelView& v = lt.view;
   {
      // cooperative interface launch: per-phase lists of the points this rank evaluates (rank-local owners)
      static_assert( sizeof( HbIfaceItem ) == 20, "HbIfaceItem layout" );
      v.ifaceFaces = v.N <= hb::tuning( "sor_iface_lists_max_n", 1024 ) ? 1 : 0;
      std::vector< std::vector< HbIfaceItem > > lists( 6 );
      const int N = v.N;
      static const int dirs[14][3] = { { -1, 0, 0 }, { 1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { -1, 1, 0 }, { 1, -1, 0 }, { 0, 0, 1 },
                                       { -1, 0, 1 }, { 0, -1, 1 }, { 1, -1, 1 }, { 0, 0, -1 }, { 0, 1, -1 }, { 1, 0, -1 }, { -1, 1, -1 } };
      // the macro primitive a point lies on, as the sorted global ids of its macro vertices packed into one key
      auto primKey = [N]( const HbCellInfo& ci, int x, int y, int z ) {
         const int b[4] = { N - x - y - z, x, y, z };
         int       g[4], n = 0;
         for ( int j = 0; j < 4; j++ )
            if ( b[j] != 0 )
               g[n++] = ci.gv[j];
         std::sort( g, g + n );
         unsigned long long key = (unsigned long long) n;
         for ( int j = 0; j < n; j++ )
            key = key * 2097152ull + (unsigned long long) ( g[j] + 1 );
         return key;
      };
      for ( int c = 0; c < view.nLocal; c++ )
      {
         const HbCellInfo& ci = view.cells[c];
         for ( const HbBndPoint& p : lt.hBnd )
         {
            const HbRegion& reg = ci.region[p.mask];
            if ( !( reg.owner & 2 ) || ( reg.dim == 2 && !v.ifaceFaces ) )
               continue;
            const int   colour = hb_interface_colour( ci, v.N, p.x, p.y, p.z );
            const int   k      = reg.dim == 0 ? 0 : ( reg.dim == 1 ? 1 + colour : 3 + colour );
            HbIfaceItem q{};
            q.cell  = c;
            q.rc0   = ci.rcBegin[p.mask];
            q.rc1   = ci.rcBegin[p.mask + 1];
            q.x     = p.x;
            q.y     = p.y;
            q.z     = p.z;
            q.bflag = (uint8_t) reg.bflag;
            // staged?  look at every neighbour of the point in every macro-cell that contains it
            const unsigned long long own = primKey( ci, p.x, p.y, p.z );
            for ( int e = q.rc0; e < q.rc1 && !q.stage; e++ )
            {
               const HbRegionCopy& rc = view.regionCopies[e];
               const HbCellInfo&   cj = view.cells[rc.slot];
               int                 P[3];
               for ( int r = 0; r < 3; r++ )
                  P[r] = rc.c[r][0] * N + rc.c[r][1] * p.x + rc.c[r][2] * p.y + rc.c[r][3] * p.z;
               for ( int d = 0; d < 14 && !q.stage; d++ )
               {
                  const int X = P[0] + dirs[d][0], Y = P[1] + dirs[d][1], Z = P[2] + dirs[d][2];
                  if ( X < 0 || Y < 0 || Z < 0 || X + Y + Z > N )
                     continue;
                  const int mq = hb_mask( N, X, Y, Z );
                  if ( mq == 0 || cj.region[mq].dim != reg.dim )
                     continue;
                  if ( hb_interface_colour( cj, N, X, Y, Z ) == colour && primKey( cj, X, Y, Z ) != own )
                     q.stage = 1;
               }
            }
            lists[k].push_back( q );
         }
      }
      for ( int k = 0; k < 6; k++ )
      {
         std::stable_sort( lists[k].begin(), lists[k].end(), []( const HbIfaceItem& a, const HbIfaceItem& b ) { return a.stage > b.stage; } );
         int n = 0;
         while ( n < (int) lists[k].size() && lists[k][n].stage )
            n++;
         v.ifaceStaged[k] = hb::tuning( "sor_stage_all", 0 ) ? (int) lists[k].size() : n;
         if ( hb::tuning( "sor_stage_all", 0 ) )
            for ( auto& it : lists[k] )
               it.stage = 1;
      }
      std::vector< HbIfaceItem > all;
      for ( int k = 0; k < 6; k++ )
      {
         v.ifaceStart[k] = (int32_t) all.size();
         all.insert( all.end(), lists[k].begin(), lists[k].end() );
      }
      v.ifaceStart[6] = (int32_t) all.size();
      lt.ifaceItems.upload( all );
      v.ifaceItems = lt.ifaceItems.ptr;
      if ( lt.scratch.n < all.size() )
         lt.scratch.alloc( all.size() );
   }
   v.phaseItems   = nullptr;
   for ( int k = 0; k <= HB_NPHASE; k++ )
      v.phaseStart[k] = 0;
   if ( !hb_sor_small_fits( v ) && v.N > 64 )
      return; // interior lists: single-CTA sweep (small levels) and whole-sweep cooperative launch (N <= 64)
   std::vector< std::vector< HbPhaseItem > > lists( HB_NPHASE );
   for ( int c = 0; c < view.nLocal; c++ )
   {
      for ( size_t l = 0; l < lt.hXyz.size(); l++ )
      {
         const HbBndPoint& q = lt.hXyz[l];
         if ( q.mask == 0 )
            lists[6 + ( ( q.x + 2 * q.y + 3 * q.z ) & 3 )].push_back( HbPhaseItem{ c, q.x, q.y, q.z, (uint8_t) view.cells[c].region[0].bflag, 0 } );
      }
   }
   std::vector< HbPhaseItem > all;
   for ( int k = 0; k < HB_NPHASE; k++ )
   {
      v.phaseStart[k] = (int32_t) all.size();
      all.insert( all.end(), lists[k].begin(), lists[k].end() );
   }
   v.phaseStart[HB_NPHASE] = (int32_t) all.size();
   lt.phaseItems.upload( all );
   v.phaseItems = lt.phaseItems.ptr;
}

hb::LevelTables& hb_storage::level( int L )
{
   if ( tablesDirty )
      uploadTopologyTables();
   auto it = levels.find( L );
   if ( it != levels.end() )
      return *it->second;
   require( L >= 0 && L <= 10, "level out of range [0,10]" );
   auto      lt = std::make_unique< hb::LevelTables >();
   const int N  = 1 << L;
   std::vector< HbBndPoint > bnd;
   std::vector< uint8_t >    mask;
   lt->hXyz.clear();
   for ( int z = 0; z <= N; z++ )
      for ( int y = 0; y <= N - z; y++ )
         for ( int x = 0; x <= N - z - y; x++ )
         {
            HbBndPoint p;
            p.x    = (uint16_t) x;
            p.y    = (uint16_t) y;
            p.z    = (uint16_t) z;
            p.mask = (uint16_t) hb_mask( N, x, y, z );
            require( (size_t) hb_index( N, x, y, z ) == lt->hXyz.size(), "internal: macro-cell index layout mismatch" );
            lt->hXyz.push_back( p );
            mask.push_back( (uint8_t) p.mask );
            if ( p.mask )
               bnd.push_back( p );
         }
   const int tileRows = hb::tuning( "tile_rows", 16 );
   const int wpb      = 8;
   std::vector< HbTile > tiles;
   for ( int z0 = 1; z0 <= N - 3; z0 += wpb )
      for ( int y0 = 1; y0 <= N - 2 - z0; y0 += tileRows )
         for ( int x0 = 1; x0 <= N - 1 - y0 - z0; x0 += 32 )
         {
            HbTile t;
            t.z0  = (int16_t) z0;
            t.y0  = (int16_t) y0;
            t.x0  = (int16_t) x0;
            t.pad = 0;
            tiles.push_back( t );
         }
   // interface points grouped by region (mask), lexicographic inside a region; chunks of <= 256 points of one region
   std::stable_sort( bnd.begin(), bnd.end(), []( const HbBndPoint& a, const HbBndPoint& b ) { return a.mask < b.mask; } );
   std::vector< HbBndChunk > chunks;
   for ( size_t i = 0; i < bnd.size(); )
   {
      size_t j = i;
      while ( j < bnd.size() && bnd[j].mask == bnd[i].mask && j - i < 256 )
         j++;
      HbBndChunk c;
      c.first = (int32_t) i;
      c.count = (int32_t) ( j - i );
      c.mask  = bnd[i].mask;
      c.pad   = 0;
      chunks.push_back( c );
      i = j;
   }
   lt->chunks.upload( chunks );
   lt->hChunks = chunks;
   lt->hBnd    = bnd;
   lt->bnd.upload( bnd );
   lt->xyz.upload( lt->hXyz );
   lt->mask.upload( mask );
   lt->tiles.upload( tiles );
   {
      std::vector< HbItem > items;
      hb_shuffle_build_items( N, items, hb::tuning( "shuffle_ys", 0 ), hb::tuning( "shuffle_order", 0 ) );
      lt->itemsShuffle.upload( items );
   }
   lt->scratch.alloc( (size_t) view.nLocal * bnd.size() );
   HbLevelView& v = lt->view;
   v.level        = L;
   v.N            = N;
   v.cellSize     = (int) lt->hXyz.size();
   v.nCells       = view.nLocal;
   v.nSlots       = (int) view.slotGlobal.size();
   v.nBnd         = (int) bnd.size();
   v.nTiles       = (int) tiles.size();
   v.tileRows     = tileRows;
   v.bnd          = lt->bnd.ptr;
   v.mask         = lt->mask.ptr;
   v.xyz          = lt->xyz.ptr;
   v.tiles        = lt->tiles.ptr;
   v.cells        = dCells.ptr;
   v.copies       = dCopies.ptr;
   v.regionCopies = dRegionCopies.ptr;
   v.chunks       = lt->chunks.ptr;
   v.nChunks      = (int) chunks.size();
   buildPhaseItems( *lt );
   auto& ref      = *lt;
   levels[L]      = std::move( lt );
   return ref;
}

// Ghost exchange tables of a level: for every plan entry the sorted union of the layers of its regions, where the layer of
// region m is the set of points of the macro-cell within stencil distance 1 of the region (the region itself included).
hb::LevelTables::Exchange& hb_storage::exchange( int L )
{
   hb::LevelTables&           lt = level( L );
   hb::LevelTables::Exchange& ex = lt.ex;
   if ( ex.built )
      return ex;
   const int N = lt.view.N;
   std::vector< std::vector< int > > layer( HB_NMASK );
   for ( const HbBndPoint& p : lt.hXyz )
   {
      if ( p.mask == 0 )
         continue;
      for ( int d = 0; d < HB_NDIR; d++ )
      {
         const int x = p.x + hb::kDir[d][0], y = p.y + hb::kDir[d][1], z = p.z + hb::kDir[d][2];
         if ( x < 0 || y < 0 || z < 0 || x + y + z > N )
            continue;
         layer[p.mask].push_back( hb_index( N, x, y, z ) );
      }
   }
   for ( auto& l : layer )
   {
      std::sort( l.begin(), l.end() );
      l.erase( std::unique( l.begin(), l.end() ), l.end() );
   }
   // class of the travelling point (exchange group, HB_NGROUP): what an SOR phase writes is one group
   auto groupOf = [&]( int slot, int i ) {
      const HbBndPoint& p = lt.hXyz[(size_t) i];
      if ( p.mask == 0 )
         return HB_NGROUP - 1;
      const int bits = ( p.mask & 1 ) + ( ( p.mask >> 1 ) & 1 ) + ( ( p.mask >> 2 ) & 1 ) + ( ( p.mask >> 3 ) & 1 );
      if ( bits >= 3 )
         return 0;
      const int colour = hb_interface_colour( view.cells[(size_t) slot], N, p.x, p.y, p.z );
      return bits == 2 ? 1 + colour : 3 + colour;
   };
   auto offsets = [&]( const std::vector< hb::LocalView::ExchangeEntry >& plan, std::vector< long long >& off, std::vector< long long >& start,
                       std::vector< std::array< int32_t, HB_NGROUP + 1 > >& grp ) {
      off.clear();
      start.assign( ex.peers.size() + 1, 0 );
      grp.assign( ex.peers.size(), {} );
      for ( size_t k = 0; k < ex.peers.size(); k++ )
      {
         start[k] = (long long) off.size();
         std::vector< std::pair< int, long long > > el; // (group, arena offset) in (plan entry, local index) order
         for ( const auto& e : plan )
         {
            if ( e.peer != ex.peers[k] )
               continue;
            std::vector< int > u;
            for ( int m = 1; m < HB_NMASK; m++ )
               if ( ( e.masks >> m ) & 1 )
                  u.insert( u.end(), layer[m].begin(), layer[m].end() );
            std::sort( u.begin(), u.end() );
            u.erase( std::unique( u.begin(), u.end() ), u.end() );
            for ( int i : u )
               el.emplace_back( groupOf( e.slot, i ), (long long) e.slot * lt.view.cellSize + i );
         }
         // both sides of a pair list the same DoFs in the same order: the plans are in ascending global cell id on both, and the group
         // of a DoF is a property of the point (the interface colour is canonical), so a stable sort by group keeps them aligned
         std::stable_sort( el.begin(), el.end(), []( const std::pair< int, long long >& a, const std::pair< int, long long >& b ) { return a.first < b.first; } );
         int32_t pos = 0;
         for ( int g = 0; g <= HB_NGROUP; g++ )
         {
            while ( pos < (int32_t) el.size() && el[(size_t) pos].first < g )
               pos++;
            grp[k][(size_t) g] = g == HB_NGROUP ? (int32_t) el.size() : pos;
         }
         for ( const auto& x : el )
            off.push_back( x.second );
      }
      start[ex.peers.size()] = (long long) off.size();
   };
   ex.peers.clear();
   for ( const auto& e : view.sendPlan )
      ex.peers.push_back( e.peer );
   for ( const auto& e : view.recvPlan )
      ex.peers.push_back( e.peer );
   std::sort( ex.peers.begin(), ex.peers.end() );
   ex.peers.erase( std::unique( ex.peers.begin(), ex.peers.end() ), ex.peers.end() );
   std::vector< long long > so, ro;
   offsets( view.sendPlan, so, ex.sendStart, ex.sendGrp );
   offsets( view.recvPlan, ro, ex.recvStart, ex.recvGrp );
   ex.sendOff.upload( so );
   ex.recvOff.upload( ro );
   ex.sendBuf.alloc( so.size() );
   ex.recvBuf.alloc( ro.size() );
   ex.built = true;
   return ex;
}

extern "C" {

const char* hb_last_error( void )
{
   return hb::g_lastError.c_str();
}

const char* hb_version( void )
{
#ifdef HB_EMUL
   return "hyteg_b200 0.1 (HOST EMULATION - tests only)";
#else
   return "hyteg_b200 0.1 (sm_100a)";
#endif
}

int hb_device_init( int device )
{
   HB_TRY
#ifndef HB_EMUL
   int n = 0;
   cudaCheck( cudaGetDeviceCount( &n ), "cudaGetDeviceCount" );
   require( n > 0, "no CUDA device: hyteg_b200 has no CPU fallback" );
#endif
   cudaCheck( cudaSetDevice( device ), "cudaSetDevice" );
   return 0;
   HB_CATCH( 1 )
}

int hb_set_stream( void* s )
{
   hb::g_stream = (cudaStream_t) s;
   return 0;
}

int hb_synchronize( void )
{
   HB_TRY
   cudaCheck( cudaStreamSynchronize( hb::g_stream ), "cudaStreamSynchronize" );
   return 0;
   HB_CATCH( 1 )
}

int hb_set_allreduce( hb_allreduce_fn fn, void* ctx )
{
   hb::g_allreduce    = fn;
   hb::g_allreduceCtx = ctx;
   return 0;
}

int hb_set_tuning( const char* key, int value )
{
   if ( std::string( key ) == "p2p_grid_waves" )
      hb_p2p_grid_waves = ( value == -2147483647 - 1 ) ? 0 : value;
   if ( std::string( key ) == "sor_coop_ctas" )
      hb_coop_ctas_per_sm = ( value == -2147483647 - 1 ) ? 0 : value;
   if ( value == -2147483647 - 1 ) // INT_MIN: back to the built-in default
      hb::g_tuning.erase( key );
   else
      hb::g_tuning[key] = value;
   return 0;
}

hb_storage* hb_storage_create( int nV, const double* xyz, int nC, const int* cells, int rank, int nRanks )
{
   return hb_storage_create_partitioned( nV, xyz, nC, cells, rank, nRanks, nullptr );
}

hb_storage* hb_storage_create_partitioned( int nV, const double* xyz, int nC, const int* cells, int rank, int nRanks, const int* cellRank )
{
   HB_TRY
   require( nRanks >= 1 && rank >= 0 && rank < nRanks, "bad rank / n_ranks" );
   auto st = std::make_unique< hb_storage >();
   st->topo.build( nV, xyz, nC, cells );
   st->topo.partition( nRanks );
   if ( cellRank )
      for ( int c = 0; c < nC; c++ )
      {
         require( cellRank[c] >= 0 && cellRank[c] < nRanks, "cell_rank entry outside [0, n_ranks)" );
         st->topo.cellRank[(size_t) c] = cellRank[c];
      }
   st->view.rank   = rank;
   st->view.nRanks = nRanks;
   st->reducePartial.alloc( HB_REDUCE_MAX_BLOCKS );
   st->reduceOut.alloc( 8 );
   cudaCheck( cudaMallocHost( (void**) &st->hReduce, 8 * sizeof( double ) ), "cudaMallocHost" );
   st->uploadTopologyTables();
   return st.release();
   HB_CATCH( nullptr )
}

/* The whole macro mesh of a partitioned storage on ONE rank: macro-cells in rank-major order (rank 0's cells in ascending id, then rank
 * 1's, ...), same vertices, same mesh boundary flags.  Coarse multigrid levels are solved redundantly on such a replica -- a few thousand
 * DoFs per level are cheaper to recompute on every rank than to keep consistent with six ghost exchanges per smoothing sweep. */
hb_storage* hb_storage_create_replica( const hb_storage* src )
{
   HB_TRY
   require( src != nullptr, "null storage" );
   const hb::Topology& t = src->topo;
   std::vector< int >  order;
   for ( int r = 0; r < src->view.nRanks; r++ )
      for ( int c = 0; c < t.nC; c++ )
         if ( t.cellRank[(size_t) c] == r )
            order.push_back( c );
   require( (int) order.size() == t.nC, "cell ranks outside [0, n_ranks)" );
   std::vector< int > cells;
   for ( int c : order )
      for ( int j = 0; j < 4; j++ )
         cells.push_back( t.cellV[(size_t) c][(size_t) j] );
   hb_storage* rep = hb_storage_create_partitioned( t.nV, t.xyz.data(), t.nC, cells.data(), 0, 1, nullptr );
   if ( !rep )
      return nullptr;
   // mesh boundary flags: primitives are identified by their (sorted) vertex ids, which the replica shares with the original
   hb::Topology&                          u = rep->topo;
   std::map< std::array< int, 2 >, int > ef;
   std::map< std::array< int, 3 >, int > ff;
   for ( int e = 0; e < t.nE; e++ )
      ef[t.edgeV[(size_t) e]] = t.eFlag[(size_t) e];
   for ( int f = 0; f < t.nF; f++ )
      ff[t.faceV[(size_t) f]] = t.fFlag[(size_t) f];
   u.vFlag = t.vFlag;
   for ( int e = 0; e < u.nE; e++ )
      u.eFlag[(size_t) e] = ef.at( u.edgeV[(size_t) e] );
   for ( int f = 0; f < u.nF; f++ )
      u.fFlag[(size_t) f] = ff.at( u.faceV[(size_t) f] );
   for ( int c = 0; c < u.nC; c++ )
      u.cFlag[(size_t) c] = t.cFlag[(size_t) order[(size_t) c]];
   rep->tablesDirty = true;
   return rep;
   HB_CATCH( nullptr )
}

static long long replicaOffsetCells( const hb_storage* part )
{
   long long n = 0;
   for ( int c = 0; c < part->topo.nC; c++ )
      if ( part->topo.cellRank[(size_t) c] < part->view.rank )
         n++;
   return n;
}

/* replica <- the owned macro-cell arrays of every rank (an all-gather, done as an all-reduce of zero-padded blocks so that the reducer a
 * host installed for dotGlobal serves it as well); interface DoFs arrive consistent because every owned copy is */
int hb_p1function_gather_to_replica( const hb_function* part, hb_function* rep, int level )
{
   HB_TRY
   require( part && rep && rep->st->view.nRanks == 1 && rep->st->topo.nC == part->st->topo.nC, "gather_to_replica: not a replica of the function's storage" );
   const size_t cs = (size_t) part->st->level( level ).view.cellSize, nLoc = (size_t) part->st->view.nLocal, nAll = (size_t) rep->st->view.nLocal;
   const size_t off = (size_t) replicaOffsetCells( part->st );
   require( off + nLoc <= nAll && nAll * cs < ( 1ull << 31 ), "gather_to_replica: level too large for a replica" );
   cudaStream_t s = hb::currentStream();
   cudaCheck( cudaMemsetAsync( rep->arena( level ), 0, nAll * cs * sizeof( double ), s ), "cudaMemsetAsync" );
   cudaCheck( cudaMemcpyAsync( rep->arena( level ) + off * cs, part->arena( level ), nLoc * cs * sizeof( double ), cudaMemcpyDeviceToDevice, s ), "cudaMemcpyAsync D2D" );
   if ( part->st->view.nRanks > 1 && !hb::deviceAllReduce( rep->arena( level ), (int) ( nAll * cs ), 0 ) )
   {
      require( hb::g_allreduce != nullptr, "gather_to_replica on a partitioned storage without a reducer: call hb_nccl_init or hb_set_allreduce first" );
      cudaCheck( cudaStreamSynchronize( s ), "cudaStreamSynchronize" );
#ifdef HB_EMUL
      hb::g_allreduce( rep->arena( level ), (int) ( nAll * cs ), 0, hb::g_allreduceCtx );
#else
      std::vector< double > h( nAll * cs );
      cudaCheck( cudaMemcpy( h.data(), rep->arena( level ), h.size() * sizeof( double ), cudaMemcpyDeviceToHost ), "cudaMemcpy D2H" );
      hb::g_allreduce( h.data(), (int) h.size(), 0, hb::g_allreduceCtx );
      cudaCheck( cudaMemcpy( rep->arena( level ), h.data(), h.size() * sizeof( double ), cudaMemcpyHostToDevice ), "cudaMemcpy H2D" );
#endif
   }
   rep->touched( level );
   return 0;
   HB_CATCH( 1 )
}

/* owned macro-cell arrays <- their block of the replica (no communication) */
int hb_p1function_scatter_from_replica( hb_function* part, const hb_function* rep, int level )
{
   HB_TRY
   require( part && rep && rep->st->view.nRanks == 1 && rep->st->topo.nC == part->st->topo.nC, "scatter_from_replica: not a replica of the function's storage" );
   const size_t cs = (size_t) part->st->level( level ).view.cellSize, nLoc = (size_t) part->st->view.nLocal, off = (size_t) replicaOffsetCells( part->st );
   cudaCheck( cudaMemcpyAsync( part->arena( level ), rep->arena( level ) + off * cs, nLoc * cs * sizeof( double ), cudaMemcpyDeviceToDevice, hb::currentStream() ),
              "cudaMemcpyAsync D2D" );
   part->touched( level );
   return 0;
   HB_CATCH( 1 )
}

void hb_p2_storage_released( hb_storage* st ); // runtime_p2.inc
void hb_storage_destroy( hb_storage* st )
{
   hb_p2_storage_released( st );
   delete st;
}

int hb_storage_set_mesh_boundary_flags_on_boundary( hb_storage* st, int fb, int fi, int highestDimAlwaysInner )
{
   HB_TRY
   require( fb >= 0 && fb < 16 && fi >= 0 && fi < 16, "mesh boundary flags must be in [0,15]" );
   st->topo.setMeshBoundaryFlagsOnBoundary( fb, fi, highestDimAlwaysInner != 0 );
   st->tablesDirty = true;
   return 0;
   HB_CATCH( 1 )
}

int hb_storage_counts( const hb_storage* st, int* c )
{
   c[0] = st->topo.nV;
   c[1] = st->topo.nE;
   c[2] = st->topo.nF;
   c[3] = st->topo.nC;
   return 0;
}

int hb_storage_primitive_vertices( const hb_storage* st, int dim, int* out )
{
   HB_TRY
   const hb::Topology& t = st->topo;
   switch ( dim )
   {
   case 0:
      for ( int i = 0; i < t.nV; i++ )
         out[i] = i;
      break;
   case 1:
      for ( int i = 0; i < t.nE; i++ )
         for ( int k = 0; k < 2; k++ )
            out[2 * i + k] = t.edgeV[i][k];
      break;
   case 2:
      for ( int i = 0; i < t.nF; i++ )
         for ( int k = 0; k < 3; k++ )
            out[3 * i + k] = t.faceV[i][k];
      break;
   case 3:
      for ( int i = 0; i < t.nC; i++ )
         for ( int k = 0; k < 4; k++ )
            out[4 * i + k] = t.cellV[i][k];
      break;
   default:
      require( false, "dim must be 0..3" );
   }
   return 0;
   HB_CATCH( 1 )
}

static std::vector< int >& flagsOf( hb::Topology& t, int dim )
{
   require( dim >= 0 && dim <= 3, "dim must be 0..3" );
   return dim == 0 ? t.vFlag : dim == 1 ? t.eFlag : dim == 2 ? t.fFlag : t.cFlag;
}

int hb_storage_get_mesh_boundary_flags( const hb_storage* st, int dim, int* out )
{
   HB_TRY
   auto& f = flagsOf( const_cast< hb::Topology& >( st->topo ), dim );
   std::memcpy( out, f.data(), f.size() * sizeof( int ) );
   return 0;
   HB_CATCH( 1 )
}

int hb_storage_set_mesh_boundary_flags( hb_storage* st, int dim, const int* in )
{
   HB_TRY
   auto& f = flagsOf( st->topo, dim );
   for ( size_t i = 0; i < f.size(); i++ )
   {
      require( in[i] >= 0 && in[i] < 16, "mesh boundary flags must be in [0,15]" );
      f[i] = in[i];
   }
   st->tablesDirty = true;
   return 0;
   HB_CATCH( 1 )
}

int hb_storage_num_local_cells( const hb_storage* st )
{
   return st->view.nLocal;
}

int hb_storage_local_cell_global_id( const hb_storage* st, int l )
{
   return ( l >= 0 && l < st->view.nLocal ) ? st->view.slotGlobal[l] : -1;
}

long long hb_cell_array_size( int level )
{
   const long long w = ( 1LL << level ) + 1;
   return w * ( w + 1 ) * ( w + 2 ) / 6;
}

long long hb_storage_num_global_dofs( const hb_storage* st, int level )
{
   const long long     N = 1LL << level;
   const hb::Topology& t = st->topo;
   const long long     e = N - 1;
   const long long     f = N >= 3 ? ( N - 1 ) * ( N - 2 ) / 2 : 0;
   const long long     c = N >= 4 ? ( N - 1 ) * ( N - 2 ) * ( N - 3 ) / 6 : 0;
   return t.nV + e * t.nE + f * t.nF + c * t.nC;
}

int hb_storage_micro_coordinates( const hb_storage* cst, int level, double* xyz )
{
   HB_TRY
   hb_storage*         st = const_cast< hb_storage* >( cst );
   hb::LevelTables&    lt = st->level( level );
   const hb::Topology& t  = st->topo;
   const double        h  = 1.0 / (double) lt.view.N;
   size_t              o  = 0;
   // Barycentric combination with the macro vertices taken in ascending GLOBAL id: every macro-cell that holds a copy of an
   // interface point evaluates the same non-zero terms in the same order, so all copies get bit-identical coordinates (and
   // interpolate( expression ) bit-identical values), whatever the local vertex order of the cell.
   for ( int l = 0; l < st->view.nLocal; l++ )
   {
      const int g = st->view.slotGlobal[l];
      int       ord[4] = { 0, 1, 2, 3 };
      std::sort( ord, ord + 4, [&]( int a, int b ) { return t.cellV[g][a] < t.cellV[g][b]; } );
      for ( const HbBndPoint& p : lt.hXyz )
      {
         const int b[4] = { lt.view.N - p.x - p.y - p.z, p.x, p.y, p.z };
         for ( int k = 0; k < 3; k++ )
         {
            double acc = 0.0;
            for ( int q = 0; q < 4; q++ )
            {
               const int j = ord[q];
               if ( b[j] != 0 )
                  acc += ( (double) b[j] * h ) * t.xyz[3 * t.cellV[g][j] + k];
            }
            xyz[o++] = acc;
         }
      }
   }
   return 0;
   HB_CATCH( 1 )
}

} // extern "C"

#include "runtime_comm.inc"
#include "runtime_function.inc"
#include "runtime_operator.inc"
#include "runtime_p2.inc"
