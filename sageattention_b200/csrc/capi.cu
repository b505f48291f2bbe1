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
extern "C" int sage_version(void) { return 100; }
