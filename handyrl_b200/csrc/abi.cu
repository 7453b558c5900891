// Error reporting and version for the C-ABI (include/b200rl.h).
#include "common.cuh"
#include "../../include/b200rl.h"

static thread_local char g_err[512] = "";

void b200rl_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* b200rl_last_error(void) { return g_err; }
extern "C" int b200rl_version(void) { return 1; }
