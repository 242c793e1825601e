// Shared host-side plumbing of the native library: error text, launch counter, device info.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>

namespace tgb {

inline char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

inline int set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return 1;
}

std::atomic<long long>& launch_counter();
inline void count_launch(int n = 1) { launch_counter().fetch_add(n, std::memory_order_relaxed); }

// Number of SMs of the current device (cached per device).
int sm_count();

}  // namespace tgb
