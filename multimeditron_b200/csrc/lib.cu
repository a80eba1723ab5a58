// lib.cu — library-level entry points and host helpers.
#include "common.cuh"

namespace mm {
thread_local int g_last_cuda_error = 0;

int sm_count() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev < 0 || dev >= 64) return 148;
  if (cached[dev].load(std::memory_order_acquire) == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev].store(n, std::memory_order_release);
  }
  return cached[dev].load(std::memory_order_acquire);
}
}  // namespace mm

extern "C" const char* mm_version(void) { return "mm_b200 0.1 (sm_100a)"; }
extern "C" int mm_last_cuda_error(void) { return mm::g_last_cuda_error; }
extern "C" int mm_device_sm_count(int device) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return n;
}
