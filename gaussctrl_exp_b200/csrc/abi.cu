// Library-level entry points of the C ABI (include/gs_b200.h).
#include <cstdlib>

#include "gs_common.cuh"

int64_t g_gs_launches = 0;

extern "C" int gs_version(void) { return 100; }
extern "C" int gs_built_arch(void) { return 100; }  // sm_100a
extern "C" int64_t gs_launch_count(void) { return g_gs_launches; }

bool gs_pdl_enabled(int group) {
    static const bool on = getenv("GS_NO_PDL") == nullptr;
    static const unsigned off_groups = [] {
        unsigned m = 0;
        if (const char *e = getenv("GS_NO_PDL_GROUPS"))
            for (; *e; ++e)
                if (*e >= '0' && *e <= '9') m |= 1u << (*e - '0');
        return m;
    }();
    return on && !((off_groups >> (group & 31)) & 1u);
}
