// Instantiations of the quant kernel for f32 inputs (one TU per dtype keeps the build parallel).
#include "quant_kernel.cuh"

namespace sglk {
int quant_launch_f32(bool fuse, int64_t group, int out_dtype, int layout, const void* x, void* q, void* s,
                      const QuantParams& p, cudaStream_t st) {
  if (fuse) return SGLK_E_DTYPE;   // the reference's v2 op has no fp32 input either
  return launch_dtype<float, false>(group, out_dtype, layout, x, q, s, p, st);
}
}  // namespace sglk
