// Error plumbing + ABI version for libclmgs.so.
#include <stdarg.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void clmgs_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

CLMGS_EXPORT const char *clmgs_last_error(void) { return g_err; }
CLMGS_EXPORT int clmgs_abi_version(void) { return 1; }
