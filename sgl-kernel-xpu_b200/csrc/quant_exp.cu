// Diagnostics: loop-structure variants of the group-quant kernel for its flagship instantiation (bf16, group 128,
// e4m3 codes, fp32 row-major scales).  Same arithmetic as ptgq_kernel (bit-identical outputs, checked by
// tools/quant_exp.py); only occupancy / persistence / prefetch structure differ.  Never used by the op.
#include "quant_kernel.cuh"

namespace sglk {

// one lane's 32-byte vector (16 bf16) of group `gid`, lane-in-group l (8 lanes per group): scale + 16 codes
template <bool kPackedAmax>
__device__ __forceinline__ void exp_process(const uint32_t (&w)[8], int64_t gid, int l, bool valid, uint8_t* __restrict__ q,
                                            float* __restrict__ s_out, float eps) {
  float2 f[8];
  float amax = 0.0f;
  if constexpr (kPackedAmax) {
    __nv_bfloat162 m = __habs2(*reinterpret_cast<const __nv_bfloat162*>(&w[0]));
#pragma unroll
    for (int i = 1; i < 8; ++i) m = __hmax2(m, __habs2(*reinterpret_cast<const __nv_bfloat162*>(&w[i])));
    const float2 mf = unpack2<__nv_bfloat16>(*reinterpret_cast<const uint32_t*>(&m));
    amax = fmaxf(mf.x, mf.y);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = unpack2<__nv_bfloat16>(w[i]);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      f[i] = unpack2<__nv_bfloat16>(w[i]);
      amax = fmaxf(fmaxf(amax, fabsf(f[i].x)), fabsf(f[i].y));
    }
  }
  amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 4));
  amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 2));
  amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 1));
  amax = fmaxf(amax, eps);
  const float rq = 1.0f / 448.0f;
  const float q0 = amax * rq;
  const float scale = __fmaf_rn(__fmaf_rn(-q0, 448.0f, amax), rq, q0);
  if (valid && l == 0) s_out[gid] = scale;
  const float r = __frcp_rn(scale);
  const float ns = -scale;
  uint32_t packed[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 c0 = div_pair(f[2 * i], r, ns), c1 = div_pair(f[2 * i + 1], r, ns);
    const uint32_t lo = __nv_cvt_float2_to_fp8x2(c0, __NV_SATFINITE, __NV_E4M3);
    const uint32_t hi = __nv_cvt_float2_to_fp8x2(c1, __NV_SATFINITE, __NV_E4M3);
    const uint32_t hib = __byte_perm(w[2 * i], w[2 * i + 1], 0x7531);
    packed[i] = ((lo | (hi << 16)) & 0x7F7F7F7Fu) | (hib & 0x80808080u);
  }
  if (valid) *reinterpret_cast<uint4*>(q + gid * 128 + l * 16) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
}

__device__ __forceinline__ void exp_load(const __nv_bfloat16* __restrict__ x, int64_t slot, int64_t total, uint32_t (&w)[8]) {
  if (slot < total) ld_stream_256(x + slot * 16, w);
  else {
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = 0;
  }
}

// U vectors per thread per chunk, all loaded before use; kPersist: grid-stride over chunks
template <int U, int MINB, bool kPersist, bool kPackedAmax>
__global__ void __launch_bounds__(256, MINB)
exp_plain_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ s, int64_t num_groups, float eps) {
  const int64_t total = num_groups * 8;
  const int tid = threadIdx.x;
  constexpr int CH = 256 * U;
  for (int64_t c0 = int64_t(blockIdx.x) * CH; c0 < total; c0 += int64_t(gridDim.x) * CH) {
    uint32_t w[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) exp_load(x, c0 + u * 256 + tid, total, w[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t slot = c0 + u * 256 + tid;
      exp_process<kPackedAmax>(w[u], slot >> 3, int(slot & 7), slot < total, q, s, eps);
    }
    if (!kPersist) break;
  }
}

// persistent, software-pipelined: the next chunk's U loads are issued before the current chunk is processed
template <int U, int MINB, bool kPackedAmax>
__global__ void __launch_bounds__(256, MINB)
exp_prefetch_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ s, int64_t num_groups, float eps) {
  const int64_t total = num_groups * 8;
  const int tid = threadIdx.x;
  constexpr int CH = 256 * U;
  const int64_t stride = int64_t(gridDim.x) * CH;
  int64_t c0 = int64_t(blockIdx.x) * CH;
  if (c0 >= total) return;
  uint32_t cur[U][8], nxt[U][8];
#pragma unroll
  for (int u = 0; u < U; ++u) exp_load(x, c0 + u * 256 + tid, total, cur[u]);
  for (;;) {
    const int64_t n0 = c0 + stride;
    const bool more = n0 < total;
    if (more) {
#pragma unroll
      for (int u = 0; u < U; ++u) exp_load(x, n0 + u * 256 + tid, total, nxt[u]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t slot = c0 + u * 256 + tid;
      exp_process<kPackedAmax>(cur[u], slot >> 3, int(slot & 7), slot < total, q, s, eps);
    }
    if (!more) break;
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i) cur[u][i] = nxt[u][i];
    c0 = n0;
  }
}

// trimmed: full-chunk fast path (no per-slot predicates, base pointers hoisted), single-LOP3 sign merge;
// kCheap: x*r only and no sign fix — NOT exact, a diagnostic for the issue-bound ceiling of this structure
template <bool kCheap>
__device__ __forceinline__ void exp_process_trim(const uint32_t (&w)[8], uint8_t* __restrict__ qdst, float* __restrict__ sdst,
                                                 bool lead, float eps) {
  float2 f[8];
  float amax = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    f[i] = unpack2<__nv_bfloat16>(w[i]);
    amax = fmaxf(fmaxf(amax, fabsf(f[i].x)), fabsf(f[i].y));
  }
  amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 4));
  amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 2));
  amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 1));
  amax = fmaxf(amax, eps);
  const float rq = 1.0f / 448.0f;
  const float q0 = amax * rq;
  const float scale = __fmaf_rn(__fmaf_rn(-q0, 448.0f, amax), rq, q0);
  if (lead) *sdst = scale;
  const float r = __frcp_rn(scale);
  const float ns = -scale;
  uint32_t packed[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 c0, c1;
    if constexpr (kCheap) {
      c0 = mul2(f[2 * i], make_float2(r, r)); c1 = mul2(f[2 * i + 1], make_float2(r, r));
    } else {
      c0 = div_pair(f[2 * i], r, ns); c1 = div_pair(f[2 * i + 1], r, ns);
    }
    const uint32_t lo = __nv_cvt_float2_to_fp8x2(c0, __NV_SATFINITE, __NV_E4M3);
    const uint32_t hi = __nv_cvt_float2_to_fp8x2(c1, __NV_SATFINITE, __NV_E4M3);
    const uint32_t mag = __byte_perm(lo, hi, 0x5410);
    if constexpr (kCheap) packed[i] = mag;
    else {
      const uint32_t hib = __byte_perm(w[2 * i], w[2 * i + 1], 0x7531);
      uint32_t o;
      asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(o) : "r"(hib), "r"(mag), "r"(0x80808080u));   // c ? a : b per bit
      packed[i] = o;
    }
  }
  *reinterpret_cast<uint4*>(qdst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
}

template <int U, int MINB, bool kCheap>
__global__ void __launch_bounds__(256, MINB)
exp_trim_kernel(const __nv_bfloat16* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ s, int64_t num_groups, float eps) {
  const int64_t total = num_groups * 8;
  const int tid = threadIdx.x;
  constexpr int CH = 256 * U;
  const int64_t c0 = int64_t(blockIdx.x) * CH;
  if (c0 + CH <= total) {
    const __nv_bfloat16* xb = x + (c0 + tid) * 16;
    uint8_t* qb = q + (c0 + tid) * 16;
    float* sb = s + ((c0 + tid) >> 3);
    uint32_t w[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) ld_stream_256(xb + u * 256 * 16, w[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) exp_process_trim<kCheap>(w[u], qb + u * 256 * 16, sb + u * 32, (tid & 7) == 0, eps);
  } else {
    uint32_t w[U][8];
#pragma unroll
    for (int u = 0; u < U; ++u) exp_load(x, c0 + u * 256 + tid, total, w[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t slot = c0 + u * 256 + tid;
      exp_process<false>(w[u], slot >> 3, int(slot & 7), slot < total, q, s, eps);
    }
  }
}

}  // namespace sglk

extern "C" __attribute__((visibility("default"))) int sglk_debug_quant_exp(const void* x, void* q, void* s, int64_t num_groups,
                                                                          float eps, int variant, int ctas_per_sm, void* stream) {
  using namespace sglk;
  if (!x || !q || !s || num_groups <= 0) return SGLK_E_ARG;
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  uint8_t* qp = static_cast<uint8_t*>(q);
  float* sp = static_cast<float*>(s);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t total = num_groups * 8;
  const int sms = device_sm_count();
  auto grid = [&](int U, bool persist) -> unsigned {
    const int64_t chunks = (total + 256 * U - 1) / (256 * U);
    if (!persist) return unsigned(chunks);
    const int64_t cap = int64_t(sms) * (ctas_per_sm > 0 ? ctas_per_sm : 16);
    return unsigned(chunks < cap ? chunks : cap);
  };
  switch (variant) {
    case 1: exp_plain_kernel<4, 3, true, false><<<grid(4, true), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;   // == product structure
    case 2: exp_plain_kernel<4, 3, false, false><<<grid(4, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 3: exp_plain_kernel<2, 4, true, false><<<grid(2, true), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 4: exp_plain_kernel<2, 4, false, false><<<grid(2, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 5: exp_plain_kernel<1, 6, false, false><<<grid(1, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 6: exp_plain_kernel<1, 8, false, false><<<grid(1, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 7: exp_prefetch_kernel<2, 3, false><<<grid(2, true), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 8: exp_prefetch_kernel<1, 5, false><<<grid(1, true), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 9: exp_plain_kernel<2, 4, false, true><<<grid(2, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 10: exp_plain_kernel<1, 6, false, true><<<grid(1, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 11: exp_prefetch_kernel<2, 3, true><<<grid(2, true), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 12: exp_plain_kernel<4, 3, false, true><<<grid(4, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 13: exp_trim_kernel<4, 3, false><<<grid(4, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 14: exp_trim_kernel<4, 3, true><<<grid(4, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 15: exp_trim_kernel<2, 4, false><<<grid(2, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    case 16: exp_trim_kernel<2, 5, true><<<grid(2, false), 256, 0, st>>>(xp, qp, sp, num_groups, eps); break;
    default: return SGLK_E_VARIANT;
  }
  return int(cudaGetLastError());
}
