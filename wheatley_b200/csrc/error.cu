#include <stdarg.h>
#include "common.cuh"

namespace b200gat {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace b200gat

extern "C" const char* b200gat_last_error(void) { return b200gat::g_err; }
extern "C" int b200gat_version(void) { return 100; }
