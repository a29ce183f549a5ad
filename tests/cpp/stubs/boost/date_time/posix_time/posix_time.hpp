// TEST SCAFFOLDING: boost::posix_time::ptime as TrackBase.h:195 declares it (rT1..rT7), for a syntax check only.
#pragma once
namespace boost { namespace posix_time { struct ptime {}; } }
