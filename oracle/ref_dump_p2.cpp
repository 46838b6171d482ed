// TEST INFRASTRUCTURE ONLY: P2 / elementwise / blending commands of the reference driver (see oracle/ref_dump.cpp).
#include "ref_dump_p2.hpp"

#include "hyteg/geometry/IcosahedralShellMap.hpp"

namespace refdump {

void setShellMap( hyteg::SetupPrimitiveStorage& setup ) { hyteg::IcosahedralShellMap::setMap( setup ); }

int mainP2( const Args& a )
{
   std::fprintf( stderr, "unknown command %s\n", a.cmd.c_str() );
   return 1;
}

} // namespace refdump
