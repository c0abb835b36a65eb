// Error plumbing and ABI version of libonepose_b200.so.
#include <stdarg.h>
#include "common.cuh"

namespace opb {
static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace opb

extern "C" const char *op_last_error(void) { return opb::g_err; }
extern "C" int op_abi_version(void) { return 1; }
