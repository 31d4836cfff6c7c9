// Host-side glue of the C ABI: version, last-error text.
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace r4 {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace r4

extern "C" int rl4co_b200_abi_version(void) { return RL4CO_B200_ABI_VERSION; }
extern "C" const char* rl4co_last_error(void) { return r4::g_err; }
