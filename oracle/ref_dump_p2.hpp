// TEST INFRASTRUCTURE ONLY: shared declarations of the reference driver (oracle/ref_dump.cpp, oracle/ref_dump_p2.cpp).
#pragma once
#include <cmath>
#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "hyteg/primitivestorage/PrimitiveStorage.hpp"
#include "hyteg/primitivestorage/SetupPrimitiveStorage.hpp"
#include "hyteg/types/PointND.hpp"

namespace refdump {

struct Args
{
   std::string                          cmd;
   std::map< std::string, std::string > kv;
   double                               num( const std::string& k, double dflt ) const;
   std::string                          str( const std::string& k, const std::string& dflt = "" ) const;
   bool                                 has( const std::string& k ) const;
};

// point cloud -> value, looked up by coordinate (spatial hash with bucket width h; h must be below the micro-mesh spacing)
struct Cloud
{
   std::unordered_map< uint64_t, std::vector< int > > buckets;
   std::vector< double >                              d;
   double                                             h;
   Cloud( const std::string& path, double bucket );
   uint64_t key( double x, double y, double z ) const;
   double   operator()( const hyteg::Point3D& p ) const;
};

std::shared_ptr< hyteg::PrimitiveStorage > makeStorage( const Args& a, bool blending );
void                                       setShellMap( hyteg::SetupPrimitiveStorage& setup ); // IcosahedralShellMap::setMap
int                                        mainP2( const Args& a );

} // namespace refdump
