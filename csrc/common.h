// Shared host-side plumbing for the b200dp native library (C ABI, loaded with
// ctypes).  Every entry point returns 0 on success and a non-zero code on
// failure; b200_last_error() returns the message for the calling thread.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdarg>
#include <cstring>

#define B200_API extern "C" __attribute__((visibility("default")))

char* b200_err_buf();  // 1 KiB thread-local buffer
inline int b200_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(b200_err_buf(), 1024, fmt, ap);
  va_end(ap);
  return code ? code : -1;
}
#define B200_CUDA(expr)                                                                     \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return b200_fail((int)_e, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),     \
                       __FILE__, __LINE__);                                                 \
  } while (0)
#define B200_CHECK(cond, ...)                                  \
  do {                                                         \
    if (!(cond)) return b200_fail(-2, __VA_ARGS__);            \
  } while (0)
#define B200_LAUNCH_CHECK() B200_CUDA(cudaGetLastError())

// Opt a kernel in to > 48 KB of dynamic shared memory.  The attribute is per DEVICE, so it is set before every
// launch (a sub-microsecond driver call) rather than once per process: a second GPU used from the same process
// would otherwise fail its launches with "invalid argument".
template <typename Kern>
inline cudaError_t set_max_smem(Kern kern, int bytes) {
  return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

int b200_num_sms();      // SM count of the current device (cached per device)
int b200_compute_sms();  // b200_num_sms() minus the SMs reserved for an overlapped collective (b200_set_reserved_sms)

// cross-file C-ABI helpers
extern "C" int b200_reduce_partials(const void* partials, int nparts, int len, void* out, float scale, void* stream);
