// C-ABI odds and ends: error buffer, version.
#include "../../include/sage_b200.h"
#include "status.h"

namespace sage {
char* last_error_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}
}  // namespace sage

extern "C" const char* sage_last_error(void) { return sage::last_error_buf(); }
extern "C" int sage_version(void) { return 200; }
// sha256 of the sources this library was built from (sageattention_b200/build.py passes it); the Python binding
// compares it with the sources it sees, so that a stale binary is noticed instead of silently used
#ifndef SAGE_SRC_HASH
#define SAGE_SRC_HASH "unknown"
#endif
extern "C" const char* sage_source_hash(void) { return SAGE_SRC_HASH; }
