// per-token-group 8-bit quantisation kernel for sm_100a (bf16/f16/f32 -> e4m3fn / int8, one scale per group)
//
// Replaces /root/reference/src/sycl/per_token_group_quant_8bit.cpp:48-220 (kernel) and the fused
// silu*mul mode of /root/reference/src/sycl/per_token_group_quant_8bit_v2.cpp:255-262.  Written from the
// op's definition (tests/test_per_token_group_quant_8bit.py:23-54); decomposition, exact-division scheme
// and scale addressing are this repo's.
//
// HBM-bound streaming kernel (2 B in + 1 B out per element + 4 B per group = 3.03125 B/elem at g=128),
// so the design goal is the fewest issued instructions per byte (round-1 profile: the first version
// was issue-bound at 72 % issue-slot utilisation and 52 % DRAM):
//   * each lane owns one 32-byte vector (16 bf16): a single 256-bit LDG (no L1 allocation, L2
//     evict-first), a single 128-bit STG of 16 codes; LPG = group_bytes/32 lanes per group (8 at
//     bf16 x 128) -> 3 xor-shuffle steps;
//   * a CTA walks contiguous chunks of 256*U vectors; the U loads of a chunk are all issued before
//     any is consumed (128 B in flight per thread);
//   * RN(x/s) without div.rn: r = RN(1/s) once per group, then per PAIR of elements three packed
//     fp32x2 ops  q0 = x*r;  e = fma(q0,-s,x);  q = fma(e,r,q0)  (Markstein correction).  Equality of
//     the e4m3 code with RN(x/s) is verified exhaustively for every (x, amax) pair of bf16 and of f16
//     on the CPU (oracle_check_division_scheme) and sampled on the device
//     (sglk_selftest_division); f32 inputs keep div.rn;
//   * the code's sign bit is taken from x itself (x/s has the sign of x; this also makes -0.0 -> 0x80
//     exact, which the fma chain alone would lose) with one PRMT + one LOP3 per four codes;
//   * e4m3 rounding is the hardware cvt.rn.satfinite.e4m3x2.f32.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sglk_fp8.h"
#include "common.cuh"

namespace sglk {

__device__ __forceinline__ uint32_t fast_div(uint32_t n, uint32_t mul, uint32_t shift) {
  return mul == 0 ? (n >> shift) : (__umulhi(n, mul) >> shift);
}

struct QuantParams {
  int64_t num_groups;       // total groups
  uint32_t groups_per_row;  // hidden / group
  uint32_t div_mul;         // n / groups_per_row == umulhi(n, mul) >> shift for n < 2^31 (mul == 0: shift only)
  uint32_t div_shift;
  int64_t scale_stride;     // column-major layouts
  int64_t in_row_stride;    // elements between input rows (fused mode: 2*hidden)
  float eps, qmin, qmax, rcp_qmax;
  int need_clamp;           // 0 when the e4m3 satfinite conversion already implies [qmin, qmax]
  // ---- v2 op only (sgl_per_token_group_quant_8bit_v2; zero / null for the v1 entry points) ----
  int v2_interleave;              // column-major scales use the v2 op's "interleaved" addressing (see col_major_slot)
  uint32_t tokens_per_expert;     // rows per expert slab (== number of tokens when there is no expert dimension)
  int64_t scale_expert_stride;    // scale elements (floats / packed int32 words) between expert slabs
  const int32_t* masked_m;        // masked MoE layout: only tokens [0, masked_m[e]) of expert e are quantised
};

// Where the scale of (row, col = group in row) goes for the column-major layouts: element index of the logical
// (orow, ocol) slot, and the expert slab offset.  v1: (row, col) itself.  v2: the reference kernel flattens
// column-major, flat = token + col * tokens_per_expert, and re-splits that index ROW-major,
// (flat / G, flat % G) (per_token_group_quant_8bit_v2.cpp:200-229) — its tests undo exactly this
// (tests/test_per_token_group_quant_8bit_v2.py:807-817).
__device__ __forceinline__ void col_major_slot(const QuantParams& p, uint32_t row, uint32_t col, uint32_t& orow,
                                               uint32_t& ocol, int64_t& slab) {
  orow = row; ocol = col; slab = 0;
  if (p.v2_interleave) {
    const uint32_t e = row / p.tokens_per_expert;
    const uint32_t flat = (row - e * p.tokens_per_expert) + col * p.tokens_per_expert;
    orow = fast_div(flat, p.div_mul, p.div_shift);
    ocol = flat - orow * p.groups_per_row;
    slab = int64_t(e) * p.scale_expert_stride;
  }
}
// packed UE8M0 word (orow, ocol/4) lives at (ocol/4)*stride + orow; the v2 op also zeroes the unused tail bytes of
// the last word of a row when G % 4 != 0 (per_token_group_quant_8bit_v2.cpp:231-248)
__device__ __forceinline__ void store_packed_ue8m0(void* s_out, const QuantParams& p, uint32_t orow, uint32_t ocol,
                                                   int64_t slab, uint8_t e8) {
  uint8_t* dst = reinterpret_cast<uint8_t*>(s_out) + (slab + int64_t(ocol >> 2) * p.scale_stride + orow) * 4 + (ocol & 3);
  *dst = e8;
  if (p.v2_interleave && ocol == p.groups_per_row - 1)
    for (uint32_t i = (ocol & 3) + 1; i < 4; ++i) dst[i - (ocol & 3)] = 0;
}
// masked MoE layout: is the token that owns group `gid` below its expert's count?
__device__ __forceinline__ bool group_active(const QuantParams& p, int64_t gid) {
  if (p.masked_m == nullptr) return true;
  const uint32_t row = fast_div(uint32_t(gid), p.div_mul, p.div_shift);
  const uint32_t e = row / p.tokens_per_expert;
  return int32_t(row - e * p.tokens_per_expert) < __ldg(p.masked_m + e);
}


// 32-byte streaming load: no L1 allocation, evict-first in L2 (the activations are read once; the
// codes written next to them should stay resident for the GEMM that follows)
__device__ __forceinline__ void ld_stream_256(const void* p, uint32_t (&v)[8]) {
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "l"(p));
}

__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\t"
      "mul.rn.f32x2 rd, ra, rb;\n\tmov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 fma2f(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %5};\n\tmov.b64 rc, {%6, %7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}

template <typename T> struct InTraits;
template <> struct InTraits<__nv_bfloat16> { static constexpr int EPV = 16; static constexpr bool k16 = true; };
template <> struct InTraits<__half>        { static constexpr int EPV = 16; static constexpr bool k16 = true; };
template <> struct InTraits<float>         { static constexpr int EPV = 8;  static constexpr bool k16 = false; };

// two packed 16-bit elements -> two floats
template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t w);
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xFFFF0000u));
}
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t w) {
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}

// silu(gate)*up evaluated the way torch evaluates `F.silu(g) * u` on a low-precision tensor: silu in
// fp32 rounded to T, then the product rounded to T (tests/test_per_token_group_quant_8bit_v2.py:420-431)
template <typename T> __device__ __forceinline__ float round_to(float x);
template <> __device__ __forceinline__ float round_to<__nv_bfloat16>(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
template <> __device__ __forceinline__ float round_to<__half>(float x) { return __half2float(__float2half_rn(x)); }
template <> __device__ __forceinline__ float round_to<float>(float x) { return x; }
template <typename T> __device__ __forceinline__ float silu_mul(float g, float u) {
  const float s = round_to<T>(__fdiv_rn(g, 1.0f + expf(-g)));
  return round_to<T>(s * u);
}

// exact ceil(log2(a)) for a normal positive float: exponent, +1 if any mantissa bit is set
__device__ __forceinline__ int ceil_log2_pos(float a) {
  const uint32_t u = __float_as_uint(a);
  return int(u >> 23) - 127 + ((u & 0x7FFFFFu) ? 1 : 0);
}

// RN(x/s) for a pair, given r = RN(1/s) and ns = -s  (see header comment)
__device__ __forceinline__ float2 div_pair(float2 x, float r, float ns) {
  const float2 r2 = make_float2(r, r);
  const float2 q0 = mul2(x, r2);
  const float2 e = fma2f(q0, make_float2(ns, ns), x);
  return fma2f(e, r2, q0);
}
__device__ __forceinline__ float div_one(float x, float r, float ns) {
  const float q0 = x * r;
  return __fmaf_rn(__fmaf_rn(q0, ns, x), r, q0);
}

template <typename T, int GROUP, bool kInt8, int kLayout, bool kFuse>
struct QuantCfg {
  static constexpr int EPV = InTraits<T>::EPV;                 // elements per 32-byte vector
  static constexpr int VPG = GROUP / EPV > 0 ? GROUP / EPV : 1;
  static constexpr int LPG = VPG < 32 ? VPG : 32;              // lanes per group
  static constexpr int VPL = VPG / LPG;                        // vectors per lane per group (2 only for f32 x 512)
  static constexpr int U = kFuse ? 2 : (VPL == 1 ? 4 : 2);     // slots per thread per chunk
  static constexpr int kThreads = 256;
  static constexpr int SLOTS_PER_CHUNK = kThreads * U;
  static_assert(GROUP % EPV == 0, "group must be a whole number of 32-byte vectors");
};

// One chunk of 256*U slots.  kFull (CTA-uniform, chosen by the caller): every slot of the chunk exists, so the
// per-slot bounds predicates compile away — the kernel is sensitive to issue pressure (tools/quant_exp.py: the trimmed
// variant 13 gains ~3 % over the predicated body, an inexact multiply-only variant ~2 % more).
// NVLS multicast stores (kSt == 1 instantiations of ptgq_chunk: the broadcast form of quant_bcast.cu): one store lands
// in the same place of every GPU's copy of a symmetric buffer, replicated by the NVSwitch
__device__ __forceinline__ void mc_store_f32(float* p, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void mc_store_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(p), "f"(__uint_as_float(a)), "f"(__uint_as_float(b)), "f"(__uint_as_float(c)), "f"(__uint_as_float(d))
               : "memory");
}

// kSt: 0 = plain stores; 1 = multimem.st to an NVLS multicast address (row-major fp32 scales, 16-byte code vectors)
template <typename T, int GROUP, bool kInt8, int kLayout, bool kFuse, bool kV2, bool kFull, int kSt = 0>
__device__ __forceinline__ void ptgq_chunk(const T* __restrict__ x, uint8_t* __restrict__ q, void* __restrict__ s_out,
                                           const QuantParams& p, const int64_t chunk0, const int64_t total_slots,
                                           const int tid) {
  using C = QuantCfg<T, GROUP, kInt8, kLayout, kFuse>;
  constexpr int EPV = C::EPV, LPG = C::LPG, VPL = C::VPL, U = C::U;
  constexpr bool k16 = InTraits<T>::k16;
  constexpr bool kUe8m0 = kLayout >= SGLK_SCALE_UE8M0_ROW_MAJOR;
  constexpr bool kColMajor = kLayout == SGLK_SCALE_F32_COL_MAJOR || kLayout == SGLK_SCALE_UE8M0_COL_PACKED;
  constexpr bool full = kFull;
  {
    uint32_t w[U][VPL][8];
    uint32_t g[kFuse ? U : 1][kFuse ? VPL : 1][8];
    // ---------------- issue every load of the chunk ----------------
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t slot = chunk0 + u * C::kThreads + tid;
      if ((full || slot < total_slots) && (!kV2 || group_active(p, slot / LPG))) {
        const int64_t gid = slot / LPG;
        const int l = int(slot % LPG);
        int64_t off;
        if constexpr (kFuse) {
          const uint32_t row = fast_div(uint32_t(gid), p.div_mul, p.div_shift);
          off = int64_t(row) * p.in_row_stride + int64_t(uint32_t(gid) - row * p.groups_per_row) * GROUP;
        } else {
          off = gid * GROUP;
        }
#pragma unroll
        for (int j = 0; j < VPL; ++j) {
          const T* src = x + off + (j * LPG + l) * EPV;
          ld_stream_256(src, w[u][j]);
          if constexpr (kFuse) ld_stream_256(src + p.in_row_stride / 2, g[u][j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < VPL; ++j)
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            w[u][j][i] = 0;
            if constexpr (kFuse) g[u][j][i] = 0;
          }
      }
    }
    // ---------------- per slot: amax -> scale -> codes ----------------
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t slot = chunk0 + u * C::kThreads + tid;
      const int64_t gid = slot / LPG;
      const bool valid = (full || slot < total_slots) && (!kV2 || group_active(p, gid));   // uniform across the group's lanes
      const int l = int(slot % LPG);
      float2 f[VPL][EPV / 2];
      float amax = 0.0f;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
#pragma unroll
        for (int i = 0; i < EPV / 2; ++i) {
          if constexpr (k16) {
            f[j][i] = unpack2<T>(w[u][j][i]);
            if constexpr (kFuse) {
              const float2 up = unpack2<T>(g[u][j][i]);
              f[j][i] = make_float2(silu_mul<T>(f[j][i].x, up.x), silu_mul<T>(f[j][i].y, up.y));
            }
          } else {
            f[j][i] = make_float2(__uint_as_float(w[u][j][2 * i]), __uint_as_float(w[u][j][2 * i + 1]));
          }
          amax = fmaxf(fmaxf(amax, fabsf(f[j][i].x)), fabsf(f[j][i].y));
        }
      }
#pragma unroll
      for (int m = LPG / 2; m >= 1; m >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, m));
      amax = fmaxf(amax, p.eps);
      // scale = RN(amax / qmax): same correction scheme with the host-rounded 1/qmax for 16-bit inputs
      float scale;
      if constexpr (k16) {
        const float q0 = amax * p.rcp_qmax;
        scale = __fmaf_rn(__fmaf_rn(-q0, p.qmax, amax), p.rcp_qmax, q0);
      } else {
        scale = __fdiv_rn(amax, p.qmax);
      }
      int exp2s = 0;
      if constexpr (kUe8m0) {
        exp2s = ceil_log2_pos(fmaxf(scale, 1e-10f));
        scale = __uint_as_float(uint32_t(exp2s + 127) << 23);
      }
      if (valid && l == 0) {
        if constexpr (kLayout == SGLK_SCALE_F32_ROW_MAJOR) {
          if constexpr (kSt == 1) mc_store_f32(reinterpret_cast<float*>(s_out) + gid, scale);
          else reinterpret_cast<float*>(s_out)[gid] = scale;
        } else if constexpr (kLayout == SGLK_SCALE_UE8M0_ROW_MAJOR) {
          reinterpret_cast<uint8_t*>(s_out)[gid] = uint8_t(exp2s + 127);
        } else {
          const uint32_t row = fast_div(uint32_t(gid), p.div_mul, p.div_shift);
          const uint32_t col = uint32_t(gid) - row * p.groups_per_row;
          if constexpr (kV2) {
            uint32_t orow, ocol;
            int64_t slab;
            col_major_slot(p, row, col, orow, ocol, slab);
            if constexpr (kLayout == SGLK_SCALE_F32_COL_MAJOR)
              reinterpret_cast<float*>(s_out)[slab + int64_t(ocol) * p.scale_stride + orow] = scale;
            else
              store_packed_ue8m0(s_out, p, orow, ocol, slab, uint8_t(exp2s + 127));
          } else if constexpr (kLayout == SGLK_SCALE_F32_COL_MAJOR) {
            reinterpret_cast<float*>(s_out)[int64_t(col) * p.scale_stride + row] = scale;
          } else {  // 4 exponent bytes per int32 word; word (row, col/4) lives at (col/4)*stride + row
            reinterpret_cast<uint8_t*>(s_out)[(int64_t(col >> 2) * p.scale_stride + row) * 4 + (col & 3)] =
                uint8_t(exp2s + 127);
          }
        }
      }
      (void)kColMajor;
      const float r = __frcp_rn(scale);
      const float ns = -scale;
#pragma unroll
      for (int j = 0; j < VPL; ++j) {
        float2 c[EPV / 2];
#pragma unroll
        for (int i = 0; i < EPV / 2; ++i) {
          if constexpr (k16) {
            c[i] = div_pair(f[j][i], r, ns);
          } else {
            c[i] = make_float2(__fdiv_rn(f[j][i].x, scale), __fdiv_rn(f[j][i].y, scale));
          }
        }
        if (kInt8 || p.need_clamp) {                   // CTA-uniform
#pragma unroll
          for (int i = 0; i < EPV / 2; ++i) {
            c[i].x = fminf(fmaxf(c[i].x, p.qmin), p.qmax);
            c[i].y = fminf(fmaxf(c[i].y, p.qmin), p.qmax);
          }
        }
        uint32_t packed[EPV / 4];
#pragma unroll
        for (int i = 0; i < EPV / 4; ++i) {
          if constexpr (kInt8) {
            packed[i] = (uint32_t(uint8_t(int8_t(__float2int_rz(c[2 * i].x)))))
                      | (uint32_t(uint8_t(int8_t(__float2int_rz(c[2 * i].y)))) << 8)
                      | (uint32_t(uint8_t(int8_t(__float2int_rz(c[2 * i + 1].x)))) << 16)
                      | (uint32_t(uint8_t(int8_t(__float2int_rz(c[2 * i + 1].y)))) << 24);
          } else {
            const uint32_t lo = __nv_cvt_float2_to_fp8x2(c[2 * i], __NV_SATFINITE, __NV_E4M3);
            const uint32_t hi = __nv_cvt_float2_to_fp8x2(c[2 * i + 1], __NV_SATFINITE, __NV_E4M3);
            uint32_t mag = lo | (hi << 16);
            if constexpr (k16 && !kFuse) {
              // sign bits straight from the four 16-bit inputs: their high bytes are bytes 1,3 of each word
              const uint32_t hib = __byte_perm(w[u][j][2 * i], w[u][j][2 * i + 1], 0x7531);
              asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(mag) : "r"(hib), "r"(mag), "r"(0x80808080u));   // sign ? hib : mag
            } else if constexpr (k16) {
              const uint32_t sg = ((__float_as_uint(f[j][2 * i].x) >> 24) & 0x80u) | ((__float_as_uint(f[j][2 * i].y) >> 16) & 0x8000u) |
                                  ((__float_as_uint(f[j][2 * i + 1].x) >> 8) & 0x800000u) | (__float_as_uint(f[j][2 * i + 1].y) & 0x80000000u);
              mag = (mag & 0x7F7F7F7Fu) | sg;
            }
            packed[i] = mag;
          }
        }
        if (valid) {
          uint8_t* dst = q + gid * GROUP + (j * LPG + l) * EPV;
          if constexpr (kSt == 1) {
            static_assert(kSt != 1 || (EPV == 16 && kLayout == SGLK_SCALE_F32_ROW_MAJOR && !kFuse && !kV2), "broadcast form");
            mc_store_v4(dst, packed[0], packed[1], packed[2], packed[3]);
          } else if constexpr (EPV == 16) *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
          else *reinterpret_cast<uint2*>(dst) = make_uint2(packed[0], packed[1]);
        }
      }
    }
  }
}

// kV2: the v2 op's column-major addressing and masked MoE layout (compiled out of the v1 instantiations: the extra
// parameter reads and the predicate cost the streaming kernel 5-8 % when they were runtime switches)
template <typename T, int GROUP, bool kInt8, int kLayout, bool kFuse, bool kV2 = false>
__global__ void __launch_bounds__(256)
ptgq_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, void* __restrict__ s_out, const QuantParams p) {
  using C = QuantCfg<T, GROUP, kInt8, kLayout, kFuse>;
  const int64_t total_slots = p.num_groups * C::LPG;   // slot = (group, lane in group); == vector index when VPL == 1
  const int tid = threadIdx.x;
  for (int64_t chunk0 = int64_t(blockIdx.x) * C::SLOTS_PER_CHUNK; chunk0 < total_slots;
       chunk0 += int64_t(gridDim.x) * C::SLOTS_PER_CHUNK) {
    if (chunk0 + C::SLOTS_PER_CHUNK <= total_slots)       // CTA-uniform
      ptgq_chunk<T, GROUP, kInt8, kLayout, kFuse, kV2, true>(x, q, s_out, p, chunk0, total_slots, tid);
    else
      ptgq_chunk<T, GROUP, kInt8, kLayout, kFuse, kV2, false>(x, q, s_out, p, chunk0, total_slots, tid);
  }
}

// Alignment-agnostic fallback (input not 32-byte aligned / codes not 16-byte aligned — e.g. an oddly
// offset view): one group per 8-lane team, scalar accesses, div.rn.  Still on the device: there is no
// CPU path.  Not a performance path.
template <typename T, bool kInt8>
__global__ void __launch_bounds__(256)
ptgq_generic_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, void* __restrict__ s_out, const QuantParams p,
                    int group, int layout, int fuse) {
  const int team = threadIdx.x >> 3, tl = threadIdx.x & 7;
  for (int64_t base = int64_t(blockIdx.x) * 32; base < p.num_groups; base += int64_t(gridDim.x) * 32) {
    const int64_t gid = base + team;
    const bool valid = gid < p.num_groups && group_active(p, gid);
    const uint32_t row = valid ? fast_div(uint32_t(gid), p.div_mul, p.div_shift) : 0;
    const uint32_t col = valid ? uint32_t(gid) - row * p.groups_per_row : 0;
    const T* src = fuse ? x + int64_t(row) * p.in_row_stride + int64_t(col) * group : x + gid * group;
    float amax = 0.0f;
    if (valid)
      for (int i = tl; i < group; i += 8) {
        float v = float(src[i]);
        if (fuse) v = silu_mul<T>(v, float(src[i + p.in_row_stride / 2]));
        amax = fmaxf(amax, fabsf(v));
      }
    amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 4));
    amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 2));
    amax = fmaxf(amax, __shfl_xor_sync(0xFFFFFFFFu, amax, 1));
    amax = fmaxf(amax, p.eps);
    float scale = __fdiv_rn(amax, p.qmax);
    int exp2s = 0;
    if (layout >= SGLK_SCALE_UE8M0_ROW_MAJOR) {
      exp2s = ceil_log2_pos(fmaxf(scale, 1e-10f));
      scale = __uint_as_float(uint32_t(exp2s + 127) << 23);
    }
    if (valid) {
      if (tl == 0) {
        if (layout == SGLK_SCALE_F32_ROW_MAJOR) reinterpret_cast<float*>(s_out)[gid] = scale;
        else if (layout == SGLK_SCALE_UE8M0_ROW_MAJOR) reinterpret_cast<uint8_t*>(s_out)[gid] = uint8_t(exp2s + 127);
        else {
          uint32_t orow, ocol;
          int64_t slab;
          col_major_slot(p, row, col, orow, ocol, slab);
          if (layout == SGLK_SCALE_F32_COL_MAJOR) reinterpret_cast<float*>(s_out)[slab + int64_t(ocol) * p.scale_stride + orow] = scale;
          else store_packed_ue8m0(s_out, p, orow, ocol, slab, uint8_t(exp2s + 127));
        }
      }
      for (int i = tl; i < group; i += 8) {
        float v = float(src[i]);
        if (fuse) v = silu_mul<T>(v, float(src[i + p.in_row_stride / 2]));
        const float c = fminf(fmaxf(__fdiv_rn(v, scale), p.qmin), p.qmax);
        if (kInt8) q[gid * group + i] = uint8_t(int8_t(__float2int_rz(c)));
        else q[gid * group + i] = uint8_t(__nv_cvt_float_to_fp8(c, __NV_SATFINITE, __NV_E4M3));
      }
    }
  }
}

template <typename T, bool kInt8>
static int launch_generic(const void* x, void* q, void* s, const QuantParams& p, int group, int layout, int fuse,
                          cudaStream_t st) {
  const int64_t need = (p.num_groups + 31) / 32;
  const int64_t cap = int64_t(device_sm_count()) * 8;
  const int blocks = int(need < cap ? (need < 1 ? 1 : need) : cap);
  ptgq_generic_kernel<T, kInt8><<<blocks, 256, 0, st>>>(static_cast<const T*>(x), static_cast<uint8_t*>(q), s, p, group, layout, fuse);
  return int(cudaGetLastError());
}

template <typename T, int GROUP, bool kInt8, int kLayout, bool kFuse, bool kV2 = false>
static int launch_one(const void* x, void* q, void* s, const QuantParams& p, cudaStream_t st) {
  using C = QuantCfg<T, GROUP, kInt8, kLayout, kFuse>;
  const int64_t slots = p.num_groups * C::LPG;
  const int64_t chunks = (slots + C::SLOTS_PER_CHUNK - 1) / C::SLOTS_PER_CHUNK;
  // one CTA per chunk (the kernel's grid-stride loop then runs once): the block scheduler staggers and balances the
  // CTAs better than a capped persistent grid did — 6.8 vs 6.44 TB/s at 32768 x 7168 (tools/quant_exp.py, variants 1/2)
  const int64_t cap = (int64_t(1) << 31) - 1;
  const int blocks = int(chunks < cap ? (chunks < 1 ? 1 : chunks) : cap);
  ptgq_kernel<T, GROUP, kInt8, kLayout, kFuse, kV2><<<blocks, C::kThreads, 0, st>>>(
      static_cast<const T*>(x), static_cast<uint8_t*>(q), s, p);
  return int(cudaGetLastError());
}

template <typename T, bool kInt8, int kLayout, bool kFuse, bool kV2 = false>
static int launch_group(int64_t group, const void* x, void* q, void* s, const QuantParams& p, cudaStream_t st) {
  switch (group) {
    case 16:  return launch_one<T, 16, kInt8, kLayout, kFuse, kV2>(x, q, s, p, st);
    case 32:  return launch_one<T, 32, kInt8, kLayout, kFuse, kV2>(x, q, s, p, st);
    case 64:  return launch_one<T, 64, kInt8, kLayout, kFuse, kV2>(x, q, s, p, st);
    case 128: return launch_one<T, 128, kInt8, kLayout, kFuse, kV2>(x, q, s, p, st);
    case 256: return launch_one<T, 256, kInt8, kLayout, kFuse, kV2>(x, q, s, p, st);
    case 512: return launch_one<T, 512, kInt8, kLayout, kFuse, kV2>(x, q, s, p, st);
    default: break;
  }
  return SGLK_E_GROUP;
}

// One translation unit per input dtype instantiates this (quant_bf16.cu, quant_f16.cu, quant_f32.cu).
template <typename T, bool kFuse>
static int launch_dtype(int64_t group, int out_dtype, int layout, const void* x, void* q, void* s,
                        const QuantParams& p, cudaStream_t st) {
  const bool aligned = (reinterpret_cast<uintptr_t>(x) & 31) == 0 && (reinterpret_cast<uintptr_t>(q) & 15) == 0 &&
                       (!kFuse || (p.in_row_stride * int64_t(sizeof(T))) % 64 == 0);
  if (!aligned) {
    if (group != 16 && group != 32 && group != 64 && group != 128 && group != 256 && group != 512) return SGLK_E_GROUP;
    return out_dtype == SGLK_INT8 ? launch_generic<T, true>(x, q, s, p, int(group), layout, kFuse, st)
                                  : launch_generic<T, false>(x, q, s, p, int(group), layout, kFuse, st);
  }
  const bool v2 = p.v2_interleave != 0 || p.masked_m != nullptr;
  if (v2) {
    // v2 column-major addressing / masked layout: e4m3 instantiations; the rare int8 form takes the generic kernel
    if (out_dtype == SGLK_INT8)
      return kFuse ? SGLK_E_DTYPE : launch_generic<T, true>(x, q, s, p, int(group), layout, kFuse, st);
    switch (layout) {
      case SGLK_SCALE_F32_COL_MAJOR: return launch_group<T, false, SGLK_SCALE_F32_COL_MAJOR, kFuse, true>(group, x, q, s, p, st);
      case SGLK_SCALE_UE8M0_COL_PACKED: return launch_group<T, false, SGLK_SCALE_UE8M0_COL_PACKED, kFuse, true>(group, x, q, s, p, st);
      default: return SGLK_E_LAYOUT;
    }
  }
  if (out_dtype == SGLK_INT8) {
    if (kFuse) return SGLK_E_DTYPE;   // the v2 fused path quantises to e4m3 only
    switch (layout) {
      case SGLK_SCALE_F32_ROW_MAJOR: return launch_group<T, true, SGLK_SCALE_F32_ROW_MAJOR, false>(group, x, q, s, p, st);
      case SGLK_SCALE_F32_COL_MAJOR: return launch_group<T, true, SGLK_SCALE_F32_COL_MAJOR, false>(group, x, q, s, p, st);
      default: return SGLK_E_LAYOUT;  // UE8M0 scales go with e4m3 codes
    }
  }
  switch (layout) {
    case SGLK_SCALE_F32_ROW_MAJOR: return launch_group<T, false, SGLK_SCALE_F32_ROW_MAJOR, kFuse>(group, x, q, s, p, st);
    case SGLK_SCALE_F32_COL_MAJOR: return launch_group<T, false, SGLK_SCALE_F32_COL_MAJOR, kFuse>(group, x, q, s, p, st);
    case SGLK_SCALE_UE8M0_ROW_MAJOR: return launch_group<T, false, SGLK_SCALE_UE8M0_ROW_MAJOR, kFuse>(group, x, q, s, p, st);
    case SGLK_SCALE_UE8M0_COL_PACKED: return launch_group<T, false, SGLK_SCALE_UE8M0_COL_PACKED, kFuse>(group, x, q, s, p, st);
    default: return SGLK_E_LAYOUT;
  }
}

// defined in quant_<dtype>.cu
int quant_launch_bf16(bool fuse, int64_t group, int out_dtype, int layout, const void* x, void* q, void* s,
                      const QuantParams& p, cudaStream_t st);
int quant_launch_f16(bool fuse, int64_t group, int out_dtype, int layout, const void* x, void* q, void* s,
                     const QuantParams& p, cudaStream_t st);
int quant_launch_f32(bool fuse, int64_t group, int out_dtype, int layout, const void* x, void* q, void* s,
                     const QuantParams& p, cudaStream_t st);

}  // namespace sglk
