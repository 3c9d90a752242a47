// Error reporting and ABI version for the C boundary.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_last_error[512] = "";

void ssar_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
    va_end(ap);
}

extern "C" const char* ssar_last_error(void) { return g_last_error; }
extern "C" int ssar_abi_version(void) { return SSAR_ABI_VERSION; }
