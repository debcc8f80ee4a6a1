// C-ABI housekeeping: error string, ABI version.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace scvi {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace scvi

extern "C" {
int scvi_b200_abi_version(void) { return 2; }
const char* scvi_last_error_string(void) { return scvi::g_err; }
}
