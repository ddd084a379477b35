#include "host.h"
#include "../../include/dtdl_b200.h"

extern "C" const char* dtdl_last_error(void) { return dtdl::get_last_error(); }
extern "C" int dtdl_version(void) { return 100; }
