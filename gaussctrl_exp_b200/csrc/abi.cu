// Library-level entry points of the C ABI (include/gs_b200.h).
#include <cstdlib>

#include "gs_common.cuh"

int64_t g_gs_launches = 0;

extern "C" int gs_version(void) { return 100; }
extern "C" int gs_built_arch(void) { return 100; }  // sm_100a
extern "C" int64_t gs_launch_count(void) { return g_gs_launches; }

bool gs_pdl_enabled() {
    static const bool on = getenv("GS_NO_PDL") == nullptr;
    return on;
}
