// per-token-group 8-bit quantisation for sm_100a  (bf16/f16/f32 -> e4m3fn / int8 + one scale per group)
//
// Replaces /root/reference/src/sycl/per_token_group_quant_8bit.cpp:48-386 (kernel + host launcher)
// and the fused silu*mul mode of /root/reference/src/sycl/per_token_group_quant_8bit_v2.cpp:255-262.
// Written from the op's definition (tests/test_per_token_group_quant_8bit.py:23-54), not from the
// SYCL text: the work decomposition, the exact-division scheme and the scale addressing are ours.
//
// HBM-bound streaming kernel: 2 B in + 1 B out per element + 4 B per group (3.03125 B/elem at g=128).
//   * one 16-byte vector per lane, LPG = group_bytes/16 lanes per group (16 for bf16 x 128), the
//     group max is an xor-butterfly over those lanes only;
//   * UNROLL independent vectors per thread are loaded before any is consumed (>= 64 B in flight
//     per thread, ~96 KB per SM at full occupancy) — enough to cover HBM latency at 6.5 TB/s;
//   * bit-exactness with the torch oracle needs RN(x/s), not x*RN(1/s).  For bf16 inputs the
//     division is done as  r = RN(1/s) (once per group);  q0 = x*r;  e = fma(-q0,s,x);
//     q = fma(e,r,q0)  — Markstein's correction, 3 FMA-pipe ops per element instead of the ~10 of
//     div.rn.  Equality of the resulting e4m3 code with RN(x/s) is checked exhaustively for every
//     (bf16 x, bf16 amax) pair on the CPU (oracle_check_division_scheme) and on the device
//     (sglk_debug_check_division).  f16/f32 inputs use __fdiv_rn.
//   * e4m3 rounding is the hardware cvt.rn.satfinite.e4m3x2.f32 (two codes per instruction).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sglk_fp8.h"
#include "common.cuh"

namespace sglk {

struct QuantParams {
  int64_t num_groups;      // total groups
  uint32_t groups_per_row; // hidden / group
  uint32_t div_mul;        // fast division by groups_per_row: q = umulhi(n, mul) >> shift (n < 2^31)
  uint32_t div_shift;
  int64_t scale_stride;    // column-major layouts
  int64_t in_row_stride;   // elements between rows of the input (fused mode: 2*hidden)
  float eps, qmin, qmax;
  int scale_layout;
  int ue8m0;
  int need_clamp;          // 0 when the e4m3 satfinite conversion already implies [qmin, qmax]
};

template <typename T> struct InTraits;
template <> struct InTraits<__nv_bfloat16> { static constexpr int EPV = 8; };
template <> struct InTraits<__half>        { static constexpr int EPV = 8; };
template <> struct InTraits<float>         { static constexpr int EPV = 4; };

template <typename T> __device__ __forceinline__ void unpack(const uint4& v, float (&f)[InTraits<T>::EPV]);
template <> __device__ __forceinline__ void unpack<__nv_bfloat16>(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
template <> __device__ __forceinline__ void unpack<__half>(const uint4& v, float (&f)[8]) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 p = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
    f[2 * i] = p.x;
    f[2 * i + 1] = p.y;
  }
}
template <> __device__ __forceinline__ void unpack<float>(const uint4& v, float (&f)[4]) {
  f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
  f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
}

// silu(gate)*up evaluated the way torch evaluates `F.silu(g) * u` on a low-precision tensor:
// silu in fp32 rounded to T, then the product rounded to T
// (tests/test_per_token_group_quant_8bit_v2.py:420-431).
template <typename T> __device__ __forceinline__ float round_to(float x);
template <> __device__ __forceinline__ float round_to<__nv_bfloat16>(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
template <> __device__ __forceinline__ float round_to<__half>(float x) { return __half2float(__float2half_rn(x)); }
template <> __device__ __forceinline__ float round_to<float>(float x) { return x; }

template <typename T> __device__ __forceinline__ float silu_mul(float g, float u) {
  const float s = round_to<T>(__fdiv_rn(g, 1.0f + expf(-g)));
  return round_to<T>(s * u);
}

__device__ __forceinline__ uint32_t fast_div(uint32_t n, uint32_t mul, uint32_t shift) {
  return mul == 0 ? (n >> shift) : (__umulhi(n, mul) >> shift);
}

// exact ceil(log2(a)) for a normal positive float: exponent, +1 if any mantissa bit is set
__device__ __forceinline__ int ceil_log2_pos(float a) {
  const uint32_t u = __float_as_uint(a);
  return int(u >> 23) - 127 + ((u & 0x7FFFFFu) ? 1 : 0);
}

template <bool kExactFast>
__device__ __forceinline__ float div_exact(float x, float s, float r) {
  if constexpr (kExactFast) {
    const float q0 = x * r;
    const float e = __fmaf_rn(-q0, s, x);
    const float q1 = __fmaf_rn(e, r, q0);
    // x = -0.0 makes e = +0 and q1 = +0; RN(x/s) keeps the sign (code 0x80), so copy it (one LOP3)
    return __uint_as_float((__float_as_uint(q1) & 0x7FFFFFFFu) | (__float_as_uint(x) & 0x80000000u));
  } else {
    return __fdiv_rn(x, s);
  }
}

template <typename T, int GROUP, bool kInt8, bool kFuseSiluMul>
__global__ void __launch_bounds__(256)
ptgq_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, void* __restrict__ s_out, const QuantParams p) {
  constexpr int EPV = InTraits<T>::EPV;            // elements per 16-byte vector
  constexpr int VPG = GROUP / EPV;                 // vectors per group
  constexpr int LPG = VPG < 32 ? VPG : 32;         // lanes per group
  constexpr int VPL = VPG / LPG;                   // vectors per lane per group
  constexpr int UNROLL = kFuseSiluMul ? 2 : (VPL == 1 ? 4 : (VPL == 2 ? 2 : 1));
  constexpr bool kFastDiv = std::is_same<T, __nv_bfloat16>::value;
  static_assert(GROUP % EPV == 0 && VPG >= 2, "group too small for 16-byte vectors");

  const int64_t total_slots = p.num_groups * LPG;  // slot = (group, lane-in-group)
  const int64_t stride = int64_t(gridDim.x) * blockDim.x;
  const int64_t first = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;

  for (int64_t base = first; base - lane < total_slots; base += stride * UNROLL) {
    uint4 v[UNROLL][VPL];
    uint4 w[kFuseSiluMul ? UNROLL : 1][kFuseSiluMul ? VPL : 1];
    // ---- issue every load of this iteration before consuming any ----
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t slot = base + int64_t(u) * stride;
      if (slot < total_slots) {
        const int64_t gid = slot / LPG;
        const int l = int(slot % LPG);
        int64_t off;
        if constexpr (kFuseSiluMul) {
          const uint32_t row = fast_div(uint32_t(gid), p.div_mul, p.div_shift);
          const uint32_t col = uint32_t(gid) - row * p.groups_per_row;
          off = int64_t(row) * p.in_row_stride + int64_t(col) * GROUP;
        } else {
          off = gid * GROUP;
        }
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
          const T* src = x + off + (j * LPG + l) * EPV;
          v[u][j] = ld_stream_v4(src);
          if constexpr (kFuseSiluMul) w[u][j] = ld_stream_v4(src + p.in_row_stride / 2);
        }
      } else {
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
          v[u][j] = make_uint4(0, 0, 0, 0);
          if constexpr (kFuseSiluMul) w[u][j] = make_uint4(0, 0, 0, 0);
        }
      }
    }
    // ---- per group: amax -> scale -> codes ----
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t slot = base + int64_t(u) * stride;
      if (slot - lane >= total_slots) break;         // warp-uniform
      const bool valid = slot < total_slots;
      float f[VPL][EPV];
      float amax = 0.0f;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        unpack<T>(v[u][j], f[j]);
        if constexpr (kFuseSiluMul) {
          float g[EPV];
          unpack<T>(w[u][j], g);
#pragma unroll
          for (int i = 0; i < EPV; ++i) f[j][i] = silu_mul<T>(f[j][i], g[i]);
        }
#pragma unroll
        for (int i = 0; i < EPV; ++i) amax = fmaxf(amax, fabsf(f[j][i]));
      }
#pragma unroll
      for (int m = LPG / 2; m >= 1; m >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, m));
      amax = fmaxf(amax, p.eps);
      float scale = __fdiv_rn(amax, p.qmax);
      int exp2s = 0;
      if (p.ue8m0) {                                  // warp-uniform
        exp2s = ceil_log2_pos(fmaxf(scale, 1e-10f));
        scale = __uint_as_float(uint32_t(exp2s + 127) << 23);
      }
      const int64_t gid = slot / LPG;
      const int l = int(slot % LPG);
      if (valid && l == 0) {
        if (p.scale_layout == SGLK_SCALE_F32_ROW_MAJOR) {
          reinterpret_cast<float*>(s_out)[gid] = scale;
        } else if (p.scale_layout == SGLK_SCALE_UE8M0_ROW_MAJOR) {
          reinterpret_cast<uint8_t*>(s_out)[gid] = uint8_t(exp2s + 127);
        } else {
          const uint32_t row = fast_div(uint32_t(gid), p.div_mul, p.div_shift);
          const uint32_t col = uint32_t(gid) - row * p.groups_per_row;
          if (p.scale_layout == SGLK_SCALE_F32_COL_MAJOR) {
            reinterpret_cast<float*>(s_out)[int64_t(col) * p.scale_stride + row] = scale;
          } else {  // 4 exponent bytes per int32 word, word (row, col/4) at (col/4)*stride + row
            reinterpret_cast<uint8_t*>(s_out)[(int64_t(col >> 2) * p.scale_stride + row) * 4 + (col & 3)] =
                uint8_t(exp2s + 127);
          }
        }
      }
      const float r = kFastDiv ? __frcp_rn(scale) : 0.0f;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        float c[EPV];
#pragma unroll
        for (int i = 0; i < EPV; ++i) c[i] = div_exact<kFastDiv>(f[j][i], scale, r);
        if (kInt8 || p.need_clamp) {                  // warp-uniform
#pragma unroll
          for (int i = 0; i < EPV; ++i) c[i] = fminf(fmaxf(c[i], p.qmin), p.qmax);
        }
        uint32_t packed[EPV / 4];
#pragma unroll
        for (int i = 0; i < EPV / 4; ++i) {
          if constexpr (kInt8) {
            packed[i] = (uint32_t(uint8_t(int8_t(__float2int_rz(c[4 * i])))))
                      | (uint32_t(uint8_t(int8_t(__float2int_rz(c[4 * i + 1])))) << 8)
                      | (uint32_t(uint8_t(int8_t(__float2int_rz(c[4 * i + 2])))) << 16)
                      | (uint32_t(uint8_t(int8_t(__float2int_rz(c[4 * i + 3])))) << 24);
          } else {
            const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(c[4 * i], c[4 * i + 1]), __NV_SATFINITE, __NV_E4M3);
            const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(c[4 * i + 2], c[4 * i + 3]), __NV_SATFINITE, __NV_E4M3);
            packed[i] = lo | (hi << 16);
          }
        }
        if (valid) {
          uint8_t* dst = q + gid * GROUP + (j * LPG + l) * EPV;
          if constexpr (EPV == 8) *reinterpret_cast<uint2*>(dst) = make_uint2(packed[0], packed[1]);
          else *reinterpret_cast<uint32_t*>(dst) = packed[0];
        }
      }
    }
  }
}

// ---- device self-check of the division scheme (used by tests, never by the op) ----
__global__ void check_division_kernel(const float* __restrict__ scales, int64_t n, unsigned long long* bad) {
  const int64_t total = n * 65536;
  unsigned long long local = 0;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += int64_t(gridDim.x) * blockDim.x) {
    const float s = scales[i >> 16];
    const float x = __uint_as_float(uint32_t(i & 0xFFFF) << 16);
    if (!(fabsf(x) <= s * 448.0f * 1.0001f)) continue;   // only |x| <= amax can occur (also drops NaN)
    const float r = __frcp_rn(s);
    const float a = div_exact<true>(x, s, r), b = __fdiv_rn(x, s);
    const uint32_t ca = __nv_cvt_float2_to_fp8x2(make_float2(a, 0.f), __NV_SATFINITE, __NV_E4M3) & 0xFF;
    const uint32_t cb = __nv_cvt_float2_to_fp8x2(make_float2(b, 0.f), __NV_SATFINITE, __NV_E4M3) & 0xFF;
    local += (ca != cb);
  }
  if (local) atomicAdd(bad, local);
}

static void make_div(uint32_t d, uint32_t* mul, uint32_t* shift) {
  // n / d for n < 2^31: power of two -> shift; else mul = ceil(2^(32+s)/d) with s = floor(log2 d)
  if ((d & (d - 1)) == 0) {
    uint32_t s = 0;
    while ((1u << s) < d) ++s;
    *mul = 0; *shift = s;
    return;
  }
  uint32_t s = 31 - __builtin_clz(d);
  const uint64_t m = ((uint64_t(1) << (32 + s)) + d - 1) / d;   // < 2^32 because d > 2^s
  *mul = uint32_t(m); *shift = s;
}

template <typename T, int GROUP, bool kInt8, bool kFuse>
static int launch_one(const void* x, void* q, void* s, const QuantParams& p, cudaStream_t st) {
  constexpr int EPV = InTraits<T>::EPV;
  constexpr int VPG = GROUP / EPV;
  constexpr int LPG = VPG < 32 ? VPG : 32;
  constexpr int VPL = VPG / LPG;
  constexpr int UNROLL = kFuse ? 2 : (VPL == 1 ? 4 : (VPL == 2 ? 2 : 1));
  const int64_t slots = p.num_groups * LPG;
  const int threads = 256;
  const int64_t need = (slots + int64_t(threads) * UNROLL - 1) / (int64_t(threads) * UNROLL);
  const int64_t cap = int64_t(device_sm_count()) * 8;     // 8 resident CTAs of 256 threads per SM
  const int blocks = int(need < cap ? (need < 1 ? 1 : need) : cap);
  ptgq_kernel<T, GROUP, kInt8, kFuse><<<blocks, threads, 0, st>>>(
      static_cast<const T*>(x), static_cast<uint8_t*>(q), s, p);
  return int(cudaGetLastError());
}

template <typename T, bool kInt8, bool kFuse>
static int launch_group(int64_t group, const void* x, void* q, void* s, const QuantParams& p, cudaStream_t st) {
  switch (group) {
    case 16:  if constexpr (InTraits<T>::EPV <= 8) return launch_one<T, 16, kInt8, kFuse>(x, q, s, p, st); break;
    case 32:  return launch_one<T, 32, kInt8, kFuse>(x, q, s, p, st);
    case 64:  return launch_one<T, 64, kInt8, kFuse>(x, q, s, p, st);
    case 128: return launch_one<T, 128, kInt8, kFuse>(x, q, s, p, st);
    case 256: return launch_one<T, 256, kInt8, kFuse>(x, q, s, p, st);
    case 512: return launch_one<T, 512, kInt8, kFuse>(x, q, s, p, st);
    default: break;
  }
  return SGLK_E_GROUP;
}

template <bool kFuse>
static int quant_dispatch(const void* input, void* output_q, void* output_s, int64_t num_groups,
                          int64_t groups_per_row, int64_t group_size, int64_t in_row_stride, int in_dtype,
                          int out_dtype, float eps, float min_8bit, float max_8bit, int scale_ue8m0,
                          int scale_layout, int64_t scale_stride, void* stream) {
  if (num_groups < 0 || groups_per_row <= 0) return SGLK_E_SHAPE;
  if (num_groups == 0) return SGLK_OK;
  if (!input || !output_q || !output_s) return SGLK_E_NULL;
  if (out_dtype != SGLK_E4M3 && out_dtype != SGLK_INT8) return SGLK_E_DTYPE;
  if (num_groups >= (int64_t(1) << 31) || num_groups % groups_per_row != 0) return SGLK_E_SHAPE;
  if (scale_layout < SGLK_SCALE_F32_ROW_MAJOR || scale_layout > SGLK_SCALE_UE8M0_COL_PACKED) return SGLK_E_LAYOUT;
  const bool layout_is_ue8m0 = scale_layout >= SGLK_SCALE_UE8M0_ROW_MAJOR;
  if (layout_is_ue8m0 != (scale_ue8m0 != 0)) return SGLK_E_LAYOUT;
  if ((scale_layout == SGLK_SCALE_F32_COL_MAJOR || scale_layout == SGLK_SCALE_UE8M0_COL_PACKED) &&
      scale_stride < num_groups / groups_per_row) return SGLK_E_STRIDE;
  if ((reinterpret_cast<uintptr_t>(input) & 15) || (reinterpret_cast<uintptr_t>(output_q) & 3)) return SGLK_E_STRIDE;
  if (!(max_8bit > 0.0f) || !(min_8bit < max_8bit)) return SGLK_E_ARG;

  QuantParams p;
  p.num_groups = num_groups;
  p.groups_per_row = uint32_t(groups_per_row);
  make_div(uint32_t(groups_per_row), &p.div_mul, &p.div_shift);
  p.scale_stride = scale_stride;
  p.in_row_stride = in_row_stride;
  p.eps = eps; p.qmin = min_8bit; p.qmax = max_8bit;
  p.scale_layout = scale_layout;
  p.ue8m0 = scale_ue8m0 ? 1 : 0;
  p.need_clamp = (out_dtype == SGLK_INT8 || min_8bit > -448.0f || max_8bit < 448.0f) ? 1 : 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool i8 = out_dtype == SGLK_INT8;
  switch (in_dtype) {
    case SGLK_BF16:
      return i8 ? launch_group<__nv_bfloat16, true, kFuse>(group_size, input, output_q, output_s, p, st)
                : launch_group<__nv_bfloat16, false, kFuse>(group_size, input, output_q, output_s, p, st);
    case SGLK_F16:
      return i8 ? launch_group<__half, true, kFuse>(group_size, input, output_q, output_s, p, st)
                : launch_group<__half, false, kFuse>(group_size, input, output_q, output_s, p, st);
    case SGLK_F32:
      if (kFuse) return SGLK_E_DTYPE;   // the reference's v2 op has no fp32 input either
      if (group_size == 16) return SGLK_E_GROUP;
      return i8 ? launch_group<float, true, false>(group_size, input, output_q, output_s, p, st)
                : launch_group<float, false, false>(group_size, input, output_q, output_s, p, st);
    default:
      return SGLK_E_DTYPE;
  }
}

}  // namespace sglk

extern "C" int sglk_per_token_group_quant_8bit(const void* input, void* output_q, void* output_s,
                                               int64_t num_groups, int64_t groups_per_row, int64_t group_size,
                                               int in_dtype, int out_dtype, float eps, float min_8bit,
                                               float max_8bit, int scale_ue8m0, int scale_layout,
                                               int64_t scale_stride, void* stream) {
  return sglk::quant_dispatch<false>(input, output_q, output_s, num_groups, groups_per_row, group_size,
                                     groups_per_row * group_size, in_dtype, out_dtype, eps, min_8bit,
                                     max_8bit, scale_ue8m0, scale_layout, scale_stride, stream);
}

extern "C" int sglk_silu_mul_per_token_group_quant_8bit(const void* input, void* output_q, void* output_s,
                                                        int64_t num_tokens, int64_t hidden, int64_t group_size,
                                                        int in_dtype, int out_dtype, float eps, float min_8bit,
                                                        float max_8bit, int scale_ue8m0, int scale_layout,
                                                        int64_t scale_stride, void* stream) {
  if (num_tokens < 0 || hidden <= 0 || group_size <= 0 || hidden % group_size != 0) return SGLK_E_SHAPE;
  const int64_t gpr = hidden / group_size;
  return sglk::quant_dispatch<true>(input, output_q, output_s, num_tokens * gpr, gpr, group_size, 2 * hidden,
                                    in_dtype, out_dtype, eps, min_8bit, max_8bit, scale_ue8m0, scale_layout,
                                    scale_stride, stream);
}

extern "C" int sglk_debug_check_division(const float* scales, int64_t num_scales,
                                         unsigned long long* mismatches, void* stream) {
  if (!scales || !mismatches || num_scales <= 0) return SGLK_E_ARG;
  sglk::check_division_kernel<<<148 * 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(scales, num_scales, mismatches);
  return int(cudaGetLastError());
}
