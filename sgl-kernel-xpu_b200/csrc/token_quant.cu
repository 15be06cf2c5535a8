// sgl_per_token_quant_fp8 for sm_100a: one fp32 scale per token (row), e4m3 codes.
//
// Replaces /root/reference/src/sycl/per_token_quant_fp8.cpp:201-306 (host launcher), :69-131 (sub-group-per-token
// kernel) and :137-199 (work-group-per-token kernel).  Arithmetic, in the reference's order (:105-123):
//     scale = rowmax|x| / 448          (IEEE division)
//     inv   = scale == 0 ? 0 : 1/scale (IEEE reciprocal)
//     q     = e4m3_rne(clamp(float(x) * inv, -448, 448))
// which is also what the reference's torch check computes from the returned scales
// (tests/test_per_token_quant_fp8.py:17-25: x.float() * scale.reciprocal(), clamp, .to(e4m3fn)).
//
// The reference reads every row twice (absmax pass, quantise pass, the second served by its L2).  Here the row
// is read from HBM ONCE and kept in registers between the two passes: a "team" of 32 (one warp) or 256 (one CTA)
// threads owns a token, each thread holds up to ITERS 16-byte (or 8-byte) vectors of it, all loads are issued
// before the first use.  Algorithmic bytes per token: hidden*(2+1) + 4.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include "../../include/sglk_fp8.h"
#include "common.cuh"

namespace sglk {

template <int BYTES> struct VecOf;
template <> struct VecOf<16> { using type = uint4; };
template <> struct VecOf<8> { using type = uint2; };

template <int BYTES>
__device__ __forceinline__ typename VecOf<BYTES>::type ld_once(const void* p) {
  return __ldcs(reinterpret_cast<const typename VecOf<BYTES>::type*>(p));   // read once: evict-first
}

// |a| max-accumulated on packed 16-bit pairs (exact: max/abs never round)
template <typename T> struct Pair;
template <> struct Pair<__nv_bfloat16> {
  using type = __nv_bfloat162;
  static __device__ __forceinline__ float2 to_float2(uint32_t w) {
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
  }
};
template <> struct Pair<__half> {
  using type = __half2;
  static __device__ __forceinline__ float2 to_float2(uint32_t w) {
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
  }
};

template <typename T>
__device__ __forceinline__ uint32_t absmax_pair(uint32_t acc, uint32_t w) {
  using P = typename Pair<T>::type;
  const P a = *reinterpret_cast<const P*>(&acc);
  const P b = __habs2(*reinterpret_cast<const P*>(&w));
  const P m = __hmax2(a, b);
  return *reinterpret_cast<const uint32_t*>(&m);
}

// two 32-bit words (4 elements) -> 4 e4m3 codes; cvt.rn.satfinite clamps to +-448 exactly like the reference's
// fmin/fmax followed by the cast (every finite |v| > 448 becomes 0x7E either way), so no separate clamp is issued
template <typename T>
__device__ __forceinline__ uint32_t quant4(uint32_t w0, uint32_t w1, float inv) {
  const float2 a = Pair<T>::to_float2(w0), b = Pair<T>::to_float2(w1);
  const uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(a.x * inv, a.y * inv), __NV_SATFINITE, __NV_E4M3);
  const uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(b.x * inv, b.y * inv), __NV_SATFINITE, __NV_E4M3);
  return lo | (hi << 16);
}

template <int TEAM>
__device__ __forceinline__ float team_max(float v, float* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
  if constexpr (TEAM > 32) {
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) smem[warp] = v;
    __syncthreads();
    v = smem[0];
#pragma unroll
    for (int i = 1; i < TEAM / 32; ++i) v = fmaxf(v, smem[i]);
  }
  return v;
}

// VB: bytes per vector (16 or 8); ITERS: vectors a thread keeps in registers (0: two passes over global memory,
// for rows too long for the register file); TEAM: threads per token (32: eight tokens per CTA; 256: one).
template <typename T, int VB, int ITERS, int TEAM>
__global__ void __launch_bounds__(256) ptq_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                  float* __restrict__ s, int64_t tokens, int64_t hidden) {
  using V = typename VecOf<VB>::type;
  constexpr int EPV = VB / 2;                 // elements per vector
  constexpr int WORDS = VB / 4;
  __shared__ float red[8];
  const int t = TEAM == 32 ? (threadIdx.x & 31) : threadIdx.x;
  const int64_t token = TEAM == 32 ? int64_t(blockIdx.x) * 8 + (threadIdx.x >> 5) : int64_t(blockIdx.x);
  if (token >= tokens) return;                // whole team leaves together (no barrier is skipped: TEAM==256 -> whole CTA)
  const int nvec = int(hidden / EPV);
  const char* row = reinterpret_cast<const char*>(x + token * hidden);
  uint8_t* qrow = q + token * hidden;

  if constexpr (ITERS > 0) {
    V v[ITERS];
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
      const int idx = t + i * TEAM;
      if (idx < nvec) v[i] = ld_once<VB>(row + size_t(idx) * VB);
      else {
        uint32_t* w = reinterpret_cast<uint32_t*>(&v[i]);
#pragma unroll
        for (int j = 0; j < WORDS; ++j) w[j] = 0u;
      }
    }
    uint32_t m = 0u;
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&v[i]);
#pragma unroll
      for (int j = 0; j < WORDS; ++j) m = absmax_pair<T>(m, w[j]);
    }
    const float2 mf = Pair<T>::to_float2(m);
    const float amax = team_max<TEAM>(fmaxf(mf.x, mf.y), red);
    const float scale = __fdiv_rn(amax, 448.0f);
    if (t == 0) s[token] = scale;
    const float inv = scale == 0.0f ? 0.0f : __frcp_rn(scale);
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
      const int idx = t + i * TEAM;
      if (idx < nvec) {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(&v[i]);
        if constexpr (VB == 16) {
          *reinterpret_cast<uint2*>(qrow + size_t(idx) * EPV) =
              make_uint2(quant4<T>(w[0], w[1], inv), quant4<T>(w[2], w[3], inv));
        } else {
          *reinterpret_cast<uint32_t*>(qrow + size_t(idx) * EPV) = quant4<T>(w[0], w[1], inv);
        }
      }
    }
  } else {
    uint32_t m = 0u;
    for (int idx = t; idx < nvec; idx += TEAM) {
      const V vv = *reinterpret_cast<const V*>(row + size_t(idx) * VB);      // default policy: re-read below from L2
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&vv);
#pragma unroll
      for (int j = 0; j < WORDS; ++j) m = absmax_pair<T>(m, w[j]);
    }
    const float2 mf = Pair<T>::to_float2(m);
    const float amax = team_max<TEAM>(fmaxf(mf.x, mf.y), red);
    const float scale = __fdiv_rn(amax, 448.0f);
    if (t == 0) s[token] = scale;
    const float inv = scale == 0.0f ? 0.0f : __frcp_rn(scale);
    for (int idx = t; idx < nvec; idx += TEAM) {
      const V vv = ld_once<VB>(row + size_t(idx) * VB);
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&vv);
      if constexpr (VB == 16) {
        *reinterpret_cast<uint2*>(qrow + size_t(idx) * EPV) =
            make_uint2(quant4<T>(w[0], w[1], inv), quant4<T>(w[2], w[3], inv));
      } else {
        *reinterpret_cast<uint32_t*>(qrow + size_t(idx) * EPV) = quant4<T>(w[0], w[1], inv);
      }
    }
  }
}

template <typename T, int VB>
static int ptq_launch(const void* x, void* q, void* s, int64_t tokens, int64_t hidden, cudaStream_t st) {
  const T* xp = static_cast<const T*>(x);
  uint8_t* qp = static_cast<uint8_t*>(q);
  float* sp = static_cast<float*>(s);
  const int64_t nvec = hidden / (VB / 2);
  // enough tokens to fill the chip with warps and a row that fits 8 vectors per lane: a warp per token
  if (tokens >= int64_t(device_sm_count()) * 16 && nvec <= 32 * 8) {
    ptq_kernel<T, VB, 8, 32><<<unsigned((tokens + 7) / 8), 256, 0, st>>>(xp, qp, sp, tokens, hidden);
    return int(cudaGetLastError());
  }
  if (tokens >= (int64_t(1) << 31)) return SGLK_E_SHAPE;
  const unsigned grid = unsigned(tokens);
  const int64_t per_thread = (nvec + 255) / 256;
  if (per_thread <= 1) ptq_kernel<T, VB, 1, 256><<<grid, 256, 0, st>>>(xp, qp, sp, tokens, hidden);
  else if (per_thread <= 2) ptq_kernel<T, VB, 2, 256><<<grid, 256, 0, st>>>(xp, qp, sp, tokens, hidden);
  else if (per_thread <= 4) ptq_kernel<T, VB, 4, 256><<<grid, 256, 0, st>>>(xp, qp, sp, tokens, hidden);
  else if (per_thread <= 8) ptq_kernel<T, VB, 8, 256><<<grid, 256, 0, st>>>(xp, qp, sp, tokens, hidden);
  else ptq_kernel<T, VB, 0, 256><<<grid, 256, 0, st>>>(xp, qp, sp, tokens, hidden);
  return int(cudaGetLastError());
}

}  // namespace sglk

extern "C" int sglk_per_token_quant_fp8(const void* input, void* output_q, void* output_s, int64_t num_tokens,
                                        int64_t hidden_dim, int in_dtype, void* stream) {
  using namespace sglk;
  if (num_tokens < 0 || hidden_dim < 0) return SGLK_E_SHAPE;
  if (hidden_dim % 4 != 0) return SGLK_E_SHAPE;                 // per_token_quant_fp8.cpp:222
  if (in_dtype != SGLK_BF16 && in_dtype != SGLK_F16) return SGLK_E_DTYPE;   // :206-208
  if (num_tokens == 0) return SGLK_OK;                           // :223
  if (!input || !output_q || !output_s) return SGLK_E_NULL;
  if (hidden_dim == 0) return SGLK_E_SHAPE;
  if (hidden_dim >= (int64_t(1) << 31)) return SGLK_E_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const uintptr_t xa = reinterpret_cast<uintptr_t>(input), qa = reinterpret_cast<uintptr_t>(output_q);
  if (reinterpret_cast<uintptr_t>(output_s) % 4 != 0) return SGLK_E_STRIDE;
  const bool wide = hidden_dim % 8 == 0 && xa % 16 == 0 && qa % 8 == 0;
  if (!wide && (xa % 8 != 0 || qa % 4 != 0)) return SGLK_E_STRIDE;
  if (in_dtype == SGLK_BF16)
    return wide ? ptq_launch<__nv_bfloat16, 16>(input, output_q, output_s, num_tokens, hidden_dim, st)
                : ptq_launch<__nv_bfloat16, 8>(input, output_q, output_s, num_tokens, hidden_dim, st);
  return wide ? ptq_launch<__half, 16>(input, output_q, output_s, num_tokens, hidden_dim, st)
              : ptq_launch<__half, 8>(input, output_q, output_s, num_tokens, hidden_dim, st);
}
