// Laboratory code, NOT part of the product library: built only into libsglk_diag.so (make diag, -DSGLK_DIAG) for tools/.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sglk_fp8.h"

// ---- diagnostics: peer (NVLink) write bandwidth vs contiguous run length --------------------------------
// Each warp store instruction writes 512 B as (512 / seg_bytes) runs of seg_bytes taken from consecutive
// rows `pitch` bytes apart — the pattern a TMA-store box with seg_bytes-wide rows produces.
__global__ void peer_write_kernel(uint8_t* __restrict__ dst, int64_t pitch, int64_t rows, int64_t width, int seg_bytes,
                                  int iters) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
  const int runs = 512 / seg_bytes;                 // rows touched per instruction
  const int64_t col_blocks = width / seg_bytes;
  const int64_t row_groups = rows / runs;
  const int64_t units = col_blocks * row_groups;
  const uint4 v = make_uint4(lane, 1, 2, 3);
  for (int it = 0; it < iters; ++it)
    for (int64_t u = warp; u < units; u += nwarps) {
      const int64_t rg = u / col_blocks, cb = u % col_blocks;
      const int64_t row = rg * runs + (lane * 16) / seg_bytes;
      const int64_t off = row * pitch + cb * seg_bytes + (lane * 16) % seg_bytes;
      *reinterpret_cast<uint4*>(dst + off) = v;
    }
}

extern "C" __attribute__((visibility("default"))) int sglk_debug_peer_write(void* dst, int64_t pitch, int64_t rows,
                                                                           int64_t width, int seg_bytes, int iters,
                                                                           void* stream) {
  if (!dst || seg_bytes < 16 || 512 % seg_bytes || width % seg_bytes || rows % (512 / seg_bytes)) return SGLK_E_ARG;
  peer_write_kernel<<<148 * 4, 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<uint8_t*>(dst), pitch, rows, width,
                                                                            seg_bytes, iters);
  return int(cudaGetLastError());
}
