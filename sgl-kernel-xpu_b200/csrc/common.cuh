// Shared device/host helpers for libsglk_sm100a.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

namespace sglk {

// 16-byte streaming load: the activations are read exactly once, keep them out of L1 and mark
// them evict-first in L2 so the codes written next to them stay resident for the GEMM.
template <typename T>
__device__ __forceinline__ uint4 ld_stream_v4(const T* p) {
  return __ldcs(reinterpret_cast<const uint4*>(p));
}

inline int device_sm_count() {
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

}  // namespace sglk
