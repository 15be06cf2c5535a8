// Instantiations of the quant kernel for f16 inputs (one TU per dtype keeps the build parallel).
#include "quant_kernel.cuh"

namespace sglk {
int quant_launch_f16(bool fuse, int64_t group, int out_dtype, int layout, const void* x, void* q, void* s,
                      const QuantParams& p, cudaStream_t st) {
  if (fuse) return launch_dtype<__half, true>(group, out_dtype, layout, x, q, s, p, st);
  return launch_dtype<__half, false>(group, out_dtype, layout, x, q, s, p, st);
}
}  // namespace sglk
