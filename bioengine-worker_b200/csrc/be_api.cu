// Library-level entry points: version and per-thread error string.
#include <stdarg.h>
#include <string.h>

#include "be_common.cuh"

static thread_local char g_err[512] = "";

void be_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" int be_version(void) { return 1000; }

extern "C" const char* be_last_error_string(void) { return g_err; }
