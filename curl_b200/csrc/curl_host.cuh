// Host-side launch helpers shared by every translation unit of libcurl_b200.so: the SM count and the
// "dynamic shared memory opt-in" of a kernel are looked up / set ONCE per device instead of on every
// call (two driver calls less per step; they matter for small images and must not sit inside a graph
// capture).  No other global state lives in the library.
#pragma once
#include <cuda_runtime.h>

namespace curl {
namespace host {

constexpr int MAX_DEVICES = 64;

static inline int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return -1;
  return dev;
}

static inline int sm_count() {
  static int cached[MAX_DEVICES] = {0};
  const int dev = current_device();
  if (dev < 0) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// One flag array per kernel: the kernel is a NON-TYPE template argument, so every instantiation gets its own
// static (a function-pointer *type* would be shared by all kernels of one signature).  `bytes` may grow
// between calls of the same kernel (segment-clamp histograms), so the largest value set so far is kept.
template <auto kernel>
static inline bool ensure_dyn_smem(size_t bytes) {
  static size_t set_to[MAX_DEVICES] = {0};
  const int dev = current_device();
  if (dev < 0) return false;
  if (set_to[dev] < bytes) {
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) return false;
    set_to[dev] = bytes;
  }
  return true;
}

// Test hook (curl_test_set_grid_cap): when > 0, no layer / colour / curve / TriSpace kernel is launched with more than
// this many blocks per image, so that SMALL test images drive every block through several loop iterations (ring-stage
// reuse, mbarrier parity flips, grid-stride steps, several tiles per block) -- the code paths the large benchmark
// shapes run.  0 (the default) leaves the launch shapes alone.  One instance per shared object: weak (every
// translation unit of the library defines it, the linker keeps one) and hidden (two copies of the library loaded
// side by side, as dev/kbench.py does, do not share it).
__attribute__((weak, visibility("hidden"))) int g_test_grid_cap = 0;

static inline int apply_grid_cap(long long n) {
  if (g_test_grid_cap > 0 && n > g_test_grid_cap) n = g_test_grid_cap;
  return (int)(n < 1 ? 1 : n);
}

}  // namespace host
}  // namespace curl
