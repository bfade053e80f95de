// Thread-local error text behind arena_last_error().
#include "host_snapshot.h"

#include <cstdarg>
#include <cstdio>

namespace {
thread_local char t_err[512] = "";
}

namespace arena_host {
int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace arena_host

extern "C" const char *arena_last_error(void) { return t_err; }
