// Shared helpers for the umetrack_b200 CUDA sources: error reporting for the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/umetrack_b200.h"

namespace umt {

char* last_error_buffer();  // thread-local, 512 bytes (defined in capi.cu)

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(UMT_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  return UMT_OK;
}

#define UMT_CUDA_TRY(expr)                                                               \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess)                                                               \
      return ::umt::fail(UMT_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e));  \
  } while (0)

#define UMT_TRY(expr)              \
  do {                             \
    int _r = (expr);               \
    if (_r != UMT_OK) return _r;   \
  } while (0)

}  // namespace umt
