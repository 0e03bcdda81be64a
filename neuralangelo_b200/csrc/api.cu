// Library-wide plumbing of libna_b200.so: ABI version and the thread-local error string.
#include "na_common.cuh"
#include <string.h>

namespace na {

static thread_local char g_error[512] = {0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

void clear_error() { g_error[0] = 0; }

}  // namespace na

extern "C" int na_abi_version(void) { return 1; }

extern "C" const char* na_last_error(void) { return na::g_error; }
