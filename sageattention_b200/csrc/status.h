// Error plumbing shared by the C-ABI entry points: thread-local message + int status.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>

namespace sage {

char* last_error_buf();  // defined in capi.cu (thread_local, 512 bytes)

inline int fail_arg(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), 512, fmt, ap);
  va_end(ap);
  return -1;
}
inline int fail_cuda(cudaError_t e, const char* what) {
  snprintf(last_error_buf(), 512, "%s: %s", what, cudaGetErrorString(e));
  return (int)e;
}

#define SAGE_REQUIRE(cond, ...) \
  do {                          \
    if (!(cond)) return ::sage::fail_arg(__VA_ARGS__); \
  } while (0)
#define SAGE_CUDA(call)                                          \
  do {                                                           \
    cudaError_t e__ = (call);                                    \
    if (e__ != cudaSuccess) return ::sage::fail_cuda(e__, #call); \
  } while (0)

}  // namespace sage
