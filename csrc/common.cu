#include "common.h"
#include <mutex>

char* b200_err_buf() {
  static thread_local char buf[1024] = {0};
  return buf;
}

B200_API const char* b200_last_error() { return b200_err_buf(); }

int b200_num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

B200_API int b200_sm_count() { return b200_num_sms(); }

// SM partition for data-parallel runs: the bucket all-reduce kernel claims `n` SMs exclusively (it asks
// for more shared memory than can co-reside with a persistent tensor-core CTA), and the persistent
// tcgen05 kernels size their grids to the remaining SMs instead of having a few of their statically
// scheduled CTAs start late behind the collective.
static int g_reserved_sms = 0;
B200_API int b200_set_reserved_sms(int n) {
  if (n < 0 || n >= b200_num_sms()) return b200_fail(-2, "reserved SM count must be in [0, SM count - 1]");
  g_reserved_sms = n;
  return 0;
}
B200_API int b200_get_reserved_sms() { return g_reserved_sms; }
int b200_compute_sms() {
  const int n = b200_num_sms() - g_reserved_sms;
  return n > 1 ? n : 1;
}

// Writes > L2-size bytes so the next timed kernel starts from a cold L2
// (B200_PROFILING.md timing hygiene).
__global__ void l2_flush_kernel(float4* buf, size_t n4, float v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
    buf[i] = make_float4(v, v, v, v);
}
B200_API int b200_l2_flush(void* buf, size_t bytes, void* stream) {
  l2_flush_kernel<<<b200_num_sms() * 8, 256, 0, (cudaStream_t)stream>>>((float4*)buf, bytes / 16, 1.0f);
  B200_LAUNCH_CHECK();
  return 0;
}
