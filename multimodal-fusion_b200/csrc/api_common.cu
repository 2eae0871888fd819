// Error state + version of the C ABI.
#include "common.cuh"

namespace hg {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
const char* get_error() { return g_err; }
}  // namespace hg

extern "C" int hg_version(void) { return 100; }   // 0.1.0
extern "C" const char* hg_last_error(void) { return hg::get_error(); }
