// Error reporting and version of the C ABI (include/egnn_b200.h).
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

static thread_local char g_err[512] = "";

void egnn_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* egnn_b200_last_error(void) { return g_err; }
extern "C" int egnn_b200_version(void) { return EGNN_B200_ABI_VERSION; }
