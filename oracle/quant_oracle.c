/*
 * CPU oracle, plain-C restatement.  TEST INFRASTRUCTURE ONLY — never linked into the product.
 *
 * Restates, with scalar IEEE-754 arithmetic only, the reference's correctness definition for the
 * FP8 block-quantized linear path:
 *   quant : /root/reference/tests/test_per_token_group_quant_8bit.py:23-40 (fp8), :43-54 (int8),
 *           :17-19 (UE8M0 rounding of the scale)
 *   e4m3  : the RNE fp32->e4m3fn rounding torch's `.to(float8_e4m3fn)` performs; the reference
 *           restates the same routine at include/sgl_kernel/jit_kernel/gemm/
 *           per_token_group_quant_8bit.hpp:42-70 — re-derived here from the format definition
 *           (1-4-3, bias 7, no inf, NaN = S.1111.111), not from that text.
 *   gemm  : /root/reference/tests/test_fp8_blockwise_gemm.py:23-63 (dequantise, fp32 product,
 *           here accumulated in double so it bounds the fp32 torch.mm from the exact side).
 *
 * Parity status: PINNED — tests/test_oracle.py checks every entry point against the golden
 * vectors produced by the reference's own python (tests/golden/make_golden.py).
 * It exists beside oracle/torch_oracle.py to prove, independently of torch's CPU kernels, that
 * "bit-exact" means: correctly rounded fp32 division twice, then RNE to e4m3.
 *
 * Build: see oracle/Makefile  (gcc -O2 -fno-fast-math -ffp-contract=off).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline float u32_as_f32(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline uint32_t f32_as_u32(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

static inline float bf16_to_f32(uint16_t h) { return u32_as_f32((uint32_t)h << 16); }

static float f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1Fu, man = h & 0x3FFu;
  if (exp == 0) {
    if (man == 0) return u32_as_f32(sign);
    float v = (float)man * 5.9604644775390625e-08f; /* 2^-24 */
    return sign ? -v : v;
  }
  if (exp == 31) return u32_as_f32(sign | 0x7F800000u | (man << 13));
  return u32_as_f32(sign | ((exp + 112u) << 23) | (man << 13));
}

/* dtype codes shared with include/sglk_fp8.h */
enum { DT_BF16 = 0, DT_F16 = 1, DT_F32 = 2 };

static inline float load_as_f32(const void* p, int64_t i, int dt) {
  switch (dt) {
    case DT_BF16: return bf16_to_f32(((const uint16_t*)p)[i]);
    case DT_F16:  return f16_to_f32(((const uint16_t*)p)[i]);
    default:      return ((const float*)p)[i];
  }
}

/* fp32 -> e4m3fn, round-to-nearest-even, by explicit rounding of the significand.
 * Values in (464, 480) round up into the S.1111.111 pattern, which *is* NaN in e4m3fn; |f| >= 480,
 * inf and NaN map to the same pattern; exactly 464 ties to even = 448 (0x7E). */
uint8_t oracle_f32_to_e4m3fn(float f) {
  uint32_t u = f32_as_u32(f);
  uint8_t sign = (uint8_t)((u >> 24) & 0x80u);
  u &= 0x7FFFFFFFu;
  if (u >= 0x43F00000u) return (uint8_t)(sign | 0x7Fu);           /* >= 480, inf, nan  */
  float a = u32_as_f32(u);
  if (a < 0.015625f) {                                            /* below 2^-6: subnormal, step 2^-9 */
    float t = a * 512.0f;                                         /* exact (power of two) */
    float r = nearbyintf(t);                                      /* RNE in default mode */
    return (uint8_t)(sign | (uint8_t)r);                          /* 0..8; 8 == 0x08 == 2^-6 */
  }
  int e = (int)(u >> 23) - 127;                                   /* -6 .. 8 */
  uint32_t man = u & 0x7FFFFFu;
  uint32_t keep = man >> 20, rest = man & 0xFFFFFu;
  if (rest > 0x80000u || (rest == 0x80000u && (keep & 1u))) keep++;
  uint32_t code = ((uint32_t)(e + 7) << 3) + keep;                /* carry propagates into exponent */
  return (uint8_t)(sign | (uint8_t)code);
}

float oracle_e4m3fn_to_f32(uint8_t c) {
  uint32_t e = (c >> 3) & 0xFu, m = c & 7u;
  float v;
  if (e == 15 && m == 7) return NAN;
  if (e == 0) v = (float)m * 0.001953125f;                        /* 2^-9 */
  else v = ldexpf(1.0f + (float)m * 0.125f, (int)e - 7);
  return (c & 0x80u) ? -v : v;
}

/* scale -> 2^ceil(log2(max(|s|,1e-10))), tests/test_per_token_group_quant_8bit.py:17-19 */
static float ceil_pow2_scale(float s, int* exp_out) {
  float a = fabsf(s);
  if (a < 1e-10f) a = 1e-10f;
  float ex = ceilf(log2f(a));
  if (exp_out) *exp_out = (int)ex;
  return exp2f(ex);
}

/*
 * x: [M,K] contiguous, K % group == 0.  q: [M,K] bytes.  s: [M,K/group] row-major fp32.
 * sbyte (optional, only with ue8m0): [M,K/group] exponent+127 bytes.
 * dst_int8 = 0 -> e4m3fn codes (no clamp, like the torch oracle); 1 -> int8 (clamp +-127, truncate).
 */
void oracle_per_token_group_quant(const void* x, int in_dtype, int64_t M, int64_t K, int64_t group,
                                  float eps, int dst_int8, int ue8m0, uint8_t* q, float* s,
                                  uint8_t* sbyte) {
  const int64_t G = K / group;
  const float qmax = dst_int8 ? 127.0f : 448.0f;
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t g = 0; g < G; ++g) {
      const int64_t base = m * K + g * group;
      float amax = 0.0f;
      int has_nan = 0;
      for (int64_t i = 0; i < group; ++i) {
        float v = fabsf(load_as_f32(x, base + i, in_dtype));
        if (v != v) has_nan = 1;
        if (v > amax) amax = v;
      }
      if (has_nan) amax = NAN;                       /* torch.amax propagates NaN */
      if (amax < eps) amax = eps;                    /* clamp(min=eps); NaN stays NaN */
      float scale = amax / qmax;                     /* correctly rounded fp32 division */
      if (ue8m0) {
        int ex;
        scale = ceil_pow2_scale(scale, &ex);
        if (sbyte) sbyte[m * G + g] = (uint8_t)(ex + 127);
      }
      s[m * G + g] = scale;
      for (int64_t i = 0; i < group; ++i) {
        float v = load_as_f32(x, base + i, in_dtype) / scale;   /* correctly rounded */
        if (dst_int8) {
          if (v < -127.0f) v = -127.0f;
          if (v > 127.0f) v = 127.0f;
          q[base + i] = (uint8_t)(int8_t)v;          /* C cast truncates toward zero, as torch */
        } else {
          q[base + i] = oracle_f32_to_e4m3fn(v);
        }
      }
    }
  }
}

/*
 * C[m,n] = sum_k (sa[m,k/128]*A[m,k]) * (sb[k/128,n/128]*B[k,n]); A [M,K] row-major codes,
 * Bt = B^T [N,K] row-major codes (the "column-major [K,N]" operand), explicit scale strides.
 * Products are formed in fp32 exactly as the torch oracle forms its operands, summed in double.
 */
void oracle_blockwise_scaled_mm(const uint8_t* A, const uint8_t* Bt, const float* sa, const float* sb,
                                int64_t M, int64_t N, int64_t K, int64_t sa_sm, int64_t sa_sk,
                                int64_t sb_sk, int64_t sb_sn, float* C) {
  float lut[256];
  for (int i = 0; i < 256; ++i) lut[i] = oracle_e4m3fn_to_f32((uint8_t)i);
  for (int64_t m = 0; m < M; ++m) {
    for (int64_t n = 0; n < N; ++n) {
      double acc = 0.0;
      for (int64_t k = 0; k < K; ++k) {
        float av = sa[m * sa_sm + (k / 128) * sa_sk] * lut[A[m * K + k]];
        float bv = sb[(k / 128) * sb_sk + (n / 128) * sb_sn] * lut[Bt[n * K + k]];
        acc += (double)av * (double)bv;
      }
      C[m * N + n] = (float)acc;
    }
  }
}

/*
 * Arithmetic check for the CUDA kernel's division scheme (DESIGN.md "quant kernel"):
 *   scale: q0 = RN(amax*rq); scale = fma(fma(-q0, qmax, amax), rq, q0)      with rq = RN(1/qmax)
 *   codes: r = RN(1/s);  q0 = RN(x*r);  e = fma(q0, -s, x);  q1 = fma(e, r, q0)
 * The scale must equal RN(amax/qmax) bit for bit, and RNE_e4m3(q1) must equal RNE_e4m3(RN(x/s)).
 * Sweeps every finite 16-bit x (bf16 when src_f16 == 0, IEEE half otherwise) with |x| <= amax against
 * every 16-bit amax whose bit pattern lies in [amax_lo, amax_hi); returns the number of mismatches.
 * (x >= 0 only: the scheme is sign-symmetric and the kernel copies the sign bit from x.)
 */
int64_t oracle_check_division_scheme(uint32_t amax_lo, uint32_t amax_hi, float qmax, float eps, int src_f16,
                                     int64_t* n_checked) {
  int64_t bad = 0, n = 0;
  const float rq = 1.0f / qmax;
  for (uint32_t ab = amax_lo; ab < amax_hi; ++ab) {
    float amax = src_f16 ? f16_to_f32((uint16_t)ab) : bf16_to_f32((uint16_t)ab);
    if (isnan(amax) || isinf(amax)) continue;
    if (amax < eps) amax = eps;                        /* clamp(min=eps) as in the quant oracle */
    float s = amax / qmax;
    float sq0 = amax * rq;
    float s2 = fmaf(fmaf(-sq0, qmax, amax), rq, sq0);
    ++n;
    if (f32_as_u32(s) != f32_as_u32(s2)) ++bad;
    float r = 1.0f / s;
    for (uint32_t xb = 0; xb <= ab; ++xb) {            /* positive x with |x| <= amax */
      float x = src_f16 ? f16_to_f32((uint16_t)xb) : bf16_to_f32((uint16_t)xb);
      float exact = x / s;
      float q0 = x * r;
      float e = fmaf(q0, -s, x);
      float q1 = fmaf(e, r, q0);
      ++n;
      if (oracle_f32_to_e4m3fn(exact) != oracle_f32_to_e4m3fn(q1)) ++bad;
    }
  }
  if (n_checked) *n_checked = n;
  return bad;
}
