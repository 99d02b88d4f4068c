#include "common.cuh"

#include <stdarg.h>
#include <string.h>

#include "../../include/dlbb.h"

namespace dlbb {

static thread_local char g_err[1024] = "";
std::atomic<uint64_t> g_launch_count{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    if (getenv("DLBB_VERBOSE_ERRORS")) fprintf(stderr, "dlbb error: %s\n", g_err);
}

}  // namespace dlbb

extern "C" {

const char* dlbb_last_error(void) { return dlbb::g_err; }

uint64_t dlbb_launch_count(void) { return dlbb::g_launch_count.load(); }

const char* dlbb_build_info(void) {
    return "dlbb sm_100a (tcgen05/TMEM/TMA), nvcc " __DATE__ " " __TIME__
#if DLBB_WATCHDOG
           " watchdog=on"
#else
           " watchdog=off"
#endif
        ;
}
}
