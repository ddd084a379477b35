#include "host.h"
#include "../../include/dtdl_b200.h"

extern "C" const char* dtdl_last_error(void) { return dtdl::get_last_error(); }
extern "C" int dtdl_version(void) { return 100; }

// NVLS probe: whether `multimem` collectives could run on this device.
// The collectives shipped in this round are the P2P variants; this is the runtime probe the
// multicast variants will be gated on.
extern "C" int dtdl_multicast_supported(int device) { return dtdl::multicast_supported(device); }
