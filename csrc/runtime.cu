// Library-wide state: launch counter, last error, device properties.
#include "tgb_internal.h"

namespace tgb {

std::atomic<long long>& launch_counter() {
  static std::atomic<long long> counter{0};
  return counter;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = n > 0 ? n : 148;
  }
  return cached[dev];
}

}  // namespace tgb

extern "C" const char* tgb_last_error() { return tgb::error_buffer(); }
extern "C" long long tgb_launch_count() { return tgb::launch_counter().load(); }
extern "C" void tgb_reset_launch_count() { tgb::launch_counter().store(0); }
