// C-ABI entry points of the per-token-group quantisation (see include/sglk_fp8.h and quant_kernel.cuh).
// Validation mirrors the reference's host launcher
// (/root/reference/src/sycl/per_token_group_quant_8bit.cpp:222-386): everything is checked before a launch.
#include "quant_kernel.cuh"

namespace sglk {

static void make_div(uint32_t d, uint32_t* mul, uint32_t* shift) {
  // n / d for n < 2^31: power of two -> shift; else mul = ceil(2^(32+s)/d) with s = floor(log2 d)
  if ((d & (d - 1)) == 0) {
    uint32_t s = 0;
    while ((1u << s) < d) ++s;
    *mul = 0; *shift = s;
    return;
  }
  const uint32_t s = 31 - __builtin_clz(d);
  const uint64_t m = ((uint64_t(1) << (32 + s)) + d - 1) / d;   // < 2^32 because d > 2^s
  *mul = uint32_t(m); *shift = s;
}

static int quant_dispatch(bool fuse, const void* input, void* output_q, void* output_s, int64_t num_groups,
                          int64_t groups_per_row, int64_t group_size, int64_t in_row_stride, int in_dtype,
                          int out_dtype, float eps, float min_8bit, float max_8bit, int scale_ue8m0,
                          int scale_layout, int64_t scale_stride, void* stream, int v2_interleave = 0,
                          int64_t tokens_per_expert = 0, int64_t scale_expert_stride = 0,
                          const int32_t* masked_m = nullptr) {
  if (num_groups < 0 || groups_per_row <= 0) return SGLK_E_SHAPE;
  if (num_groups == 0) return SGLK_OK;
  if (!input || !output_q || !output_s) return SGLK_E_NULL;
  if (in_dtype != SGLK_BF16 && in_dtype != SGLK_F16 && in_dtype != SGLK_F32) return SGLK_E_DTYPE;
  if (out_dtype != SGLK_E4M3 && out_dtype != SGLK_INT8) return SGLK_E_DTYPE;
  if (group_size != 16 && group_size != 32 && group_size != 64 && group_size != 128 && group_size != 256 &&
      group_size != 512) return SGLK_E_GROUP;
  if (num_groups >= (int64_t(1) << 31) || num_groups % groups_per_row != 0) return SGLK_E_SHAPE;
  if (scale_layout < SGLK_SCALE_F32_ROW_MAJOR || scale_layout > SGLK_SCALE_UE8M0_COL_PACKED) return SGLK_E_LAYOUT;
  const bool layout_is_ue8m0 = scale_layout >= SGLK_SCALE_UE8M0_ROW_MAJOR;
  if (layout_is_ue8m0 != (scale_ue8m0 != 0)) return SGLK_E_LAYOUT;
  if ((scale_layout == SGLK_SCALE_F32_COL_MAJOR || scale_layout == SGLK_SCALE_UE8M0_COL_PACKED) &&
      scale_stride < (tokens_per_expert > 0 ? tokens_per_expert : num_groups / groups_per_row)) return SGLK_E_STRIDE;
  const int esz = in_dtype == SGLK_F32 ? 4 : 2;
  if (reinterpret_cast<uintptr_t>(input) % esz != 0) return SGLK_E_STRIDE;
  if (!(max_8bit > 0.0f) || !(min_8bit < max_8bit)) return SGLK_E_ARG;
  if (fuse && in_dtype == SGLK_F32) return SGLK_E_DTYPE;

  QuantParams p;
  p.num_groups = num_groups;
  p.groups_per_row = uint32_t(groups_per_row);
  make_div(uint32_t(groups_per_row), &p.div_mul, &p.div_shift);
  p.scale_stride = scale_stride;
  p.in_row_stride = in_row_stride;
  p.eps = eps; p.qmin = min_8bit; p.qmax = max_8bit;
  p.rcp_qmax = 1.0f / max_8bit;                     // correctly rounded on the host
  p.need_clamp = (out_dtype == SGLK_INT8 || min_8bit > -448.0f || max_8bit < 448.0f) ? 1 : 0;
  p.v2_interleave = v2_interleave;
  p.tokens_per_expert = uint32_t(tokens_per_expert > 0 ? tokens_per_expert : num_groups / groups_per_row);
  p.scale_expert_stride = scale_expert_stride;
  p.masked_m = masked_m;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (in_dtype) {
    case SGLK_BF16: return quant_launch_bf16(fuse, group_size, out_dtype, scale_layout, input, output_q, output_s, p, st);
    case SGLK_F16:  return quant_launch_f16(fuse, group_size, out_dtype, scale_layout, input, output_q, output_s, p, st);
    default:        return quant_launch_f32(fuse, group_size, out_dtype, scale_layout, input, output_q, output_s, p, st);
  }
}

// ---- device self-check of the division scheme (used by tests, never by the op) ----
// numerators: every 16-bit pattern, read as bf16 (mode 0) or f16 (mode 1)
__global__ void check_division_kernel(const float* __restrict__ scales, int64_t n, int mode, unsigned long long* bad) {
  const int64_t total = n * 65536;
  unsigned long long local = 0;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += int64_t(gridDim.x) * blockDim.x) {
    const float s = scales[i >> 16];
    const uint32_t bits = uint32_t(i & 0xFFFF);
    const float x = mode == 0 ? __uint_as_float(bits << 16) : __half2float(__ushort_as_half((unsigned short)bits));
    if (!(fabsf(x) <= s * 448.0f * 1.0001f)) continue;   // only |x| <= amax can occur (also drops NaN)
    const float r = __frcp_rn(s);
    const float2 a2 = div_pair(make_float2(x, -x), r, -s);
    const float a = div_one(x, r, -s), b = __fdiv_rn(x, s);
    const uint32_t ca = __nv_cvt_float2_to_fp8x2(make_float2(a, a2.x), __NV_SATFINITE, __NV_E4M3);
    const uint32_t cb = __nv_cvt_float2_to_fp8x2(make_float2(b, b), __NV_SATFINITE, __NV_E4M3);
    local += ((ca & 0x7F7F) != (cb & 0x7F7F));           // magnitudes; the sign bit is copied from x by the op
  }
  if (local) atomicAdd(bad, local);
}

}  // namespace sglk

extern "C" int sglk_per_token_group_quant_8bit(const void* input, void* output_q, void* output_s,
                                               int64_t num_groups, int64_t groups_per_row, int64_t group_size,
                                               int in_dtype, int out_dtype, float eps, float min_8bit,
                                               float max_8bit, int scale_ue8m0, int scale_layout,
                                               int64_t scale_stride, void* stream) {
  return sglk::quant_dispatch(false, input, output_q, output_s, num_groups, groups_per_row, group_size,
                              groups_per_row * group_size, in_dtype, out_dtype, eps, min_8bit, max_8bit,
                              scale_ue8m0, scale_layout, scale_stride, stream);
}

extern "C" int sglk_silu_mul_per_token_group_quant_8bit(const void* input, void* output_q, void* output_s,
                                                        int64_t num_tokens, int64_t hidden, int64_t group_size,
                                                        int in_dtype, int out_dtype, float eps, float min_8bit,
                                                        float max_8bit, int scale_ue8m0, int scale_layout,
                                                        int64_t scale_stride, void* stream) {
  if (num_tokens < 0 || hidden <= 0 || group_size <= 0 || hidden % group_size != 0) return SGLK_E_SHAPE;
  const int64_t gpr = hidden / group_size;
  return sglk::quant_dispatch(true, input, output_q, output_s, num_tokens * gpr, gpr, group_size, 2 * hidden,
                              in_dtype, out_dtype, eps, min_8bit, max_8bit, scale_ue8m0, scale_layout,
                              scale_stride, stream);
}

extern "C" int sglk_per_token_group_quant_8bit_v2(const void* input, void* output_q, void* output_s,
                                                  int64_t num_experts, int64_t tokens_per_expert, int64_t hidden,
                                                  int64_t group_size, int in_dtype, int out_dtype, float eps,
                                                  float min_8bit, float max_8bit, int scale_ue8m0, int scale_layout,
                                                  int64_t scale_stride, int64_t scale_expert_stride,
                                                  int fuse_silu_and_mul, const int32_t* masked_m, void* stream) {
  if (num_experts < 0 || tokens_per_expert < 0 || hidden <= 0 || group_size <= 0 || hidden % group_size != 0)
    return SGLK_E_SHAPE;
  if (num_experts * tokens_per_expert == 0) return SGLK_OK;
  const int64_t gpr = hidden / group_size;
  if (tokens_per_expert * gpr >= (int64_t(1) << 31)) return SGLK_E_SHAPE;        // the interleave index is 32-bit
  const bool col = scale_layout == SGLK_SCALE_F32_COL_MAJOR || scale_layout == SGLK_SCALE_UE8M0_COL_PACKED;
  if (masked_m != nullptr) {
    // the reference dispatches its masked scheduler only for this combination (…_v2.cpp:779-783)
    if (!(col && scale_ue8m0 && fuse_silu_and_mul)) return SGLK_E_LAYOUT;
    if (scale_expert_stride < 0) return SGLK_E_STRIDE;
  } else if (num_experts != 1) {
    return SGLK_E_SHAPE;
  }
  return sglk::quant_dispatch(fuse_silu_and_mul != 0, input, output_q, output_s, num_experts * tokens_per_expert * gpr,
                              gpr, group_size, (fuse_silu_and_mul ? 2 : 1) * hidden, in_dtype, out_dtype, eps, min_8bit,
                              max_8bit, scale_ue8m0, scale_layout, scale_stride, stream, col ? 1 : 0,
                              tokens_per_expert, scale_expert_stride, masked_m);
}

extern "C" int sglk_selftest_division(const float* scales, int64_t num_scales, int numerator_dtype,
                                         unsigned long long* mismatches, void* stream) {
  if (!scales || !mismatches || num_scales <= 0 || (numerator_dtype != SGLK_BF16 && numerator_dtype != SGLK_F16))
    return SGLK_E_ARG;
  sglk::check_division_kernel<<<148 * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      scales, num_scales, numerator_dtype == SGLK_F16 ? 1 : 0, mismatches);
  return int(cudaGetLastError());
}
