// Host-side ABI plumbing: version + thread-local error string.
#include <stdarg.h>

#include "common.cuh"

namespace rl4co {
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace rl4co

extern "C" int rl4co_abi_version(void) { return RL4CO_ABI_VERSION; }
extern "C" const char *rl4co_last_error(void) { return rl4co::g_err; }
