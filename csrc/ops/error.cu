#include "common.cuh"

namespace pdb {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace pdb

extern "C" const char* pdb_last_error(void) { return pdb::g_err; }
extern "C" int pdb_version(void) { return 1; }
