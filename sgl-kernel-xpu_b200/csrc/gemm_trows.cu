// Row-balanced fp8 blockwise-scaled GEMM for 65..128 tokens, token-major (variant `rows_tokmajor`).
//
//   C[m, n] = sum_kb sa[m, kb] * sb[kb, n / 128] * sum_{k in kb} A[m, k] * B[k, n]
//
// Same op as gemm.cu / gemm_decode.cu (declared /root/reference/include/sgl_kernel_ops.h:581-586, semantics
// tests/test_fp8_blockwise_gemm.py:23-63; the reference ships no kernel).  At 65..128 tokens the op is still a weight
// stream (N*K bytes once), but neither decode kernel fits it: the 128x128 token-major TILES leave a second, partly filled
// wave or a stream-K fix-up (192 tiles on 148 SMs: 41 us at N=24576, 0.68 of the copy bandwidth), and the swap-AB row
// kernel pays one latency-bound MMA chain per 128 weight rows with the tokens on the narrow side (50 us).  Here, as in the
// swap-AB row kernel, every CTA owns ONE contiguous range of R = ceil8(N / #SMs) weight rows for all of K — no second
// wave, no partials, no workspace — but the tokens stay on the MMA's M side (all 128 lanes live) and the R weight rows are
// its N side: four 128 x ceil16(R) x 32 MMAs per K block whatever R is.
//   * a range is walked in passes of <= 256 rows (one accumulator stage = 256 TMEM columns, two stages);
//   * per K block: one TMA box of the pass's weight rows, one of the tokens that exist (the box is cut to ceil8(M)
//     rows; the MMA reads 128 smem rows regardless, stale rows only feed TMEM lanes that are never stored);
//   * promotion in registers as in the prefill kernel: thread = token row, warp = (lane quarter, 128-column half);
//     sa[token, kb] and the (up to three) sb[kb, .] values of the pass travel global -> smem with cp.async behind an
//     mbarrier; a 128-row scale-block boundary can fall inside a 32-column chunk, but only at multiples of 8 (ranges
//     start at multiples of 8), so the factor is formed per 8 columns;
//   * C leaves the registers directly: a thread holds one token row x up to 128 consecutive columns -> 16-byte stores.
#include "gemm_common.cuh"

namespace sglk {

constexpr int kTrMaxStages = 8;
constexpr int kTrThreads = 128 + 8 * 32;
constexpr int kTrTailBytes = 8192;            // behind the ring: factor slots + barriers

struct TRowsParams {
  const float* sa;
  const float* sb;
  void* C;
  int64_t M, N, K, ldc;
  int64_t sa_sm, sa_sk, sb_sk, sb_sn;
  int32_t num_k_blocks, num_sb_cols;
  int32_t rows_per_worker, passes, rows_per_pass;   // multiples of 8; rows_per_pass <= 256
  int32_t num_stages;
  uint32_t stage_bytes, x_off;                      // stage = [weight rows of the pass][token tile], each 1 KB aligned
  uint32_t stage_tx_bytes;
  uint32_t idesc;                                   // instruction descriptor for 128 x ceil16(rows_per_pass)
};

constexpr int kTrScaleSlot = (128 + 4) * 4;         // sa[0..128) + four sb values per K block

template <typename OutT>
__global__ void __launch_bounds__(kTrThreads, 1)
fp8_blockwise_gemm_trows_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                                const TRowsParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  if ((ptx::smem_u32(smem_raw) & 1023u) != 0) __trap();
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  const int nstages = p.num_stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t scale_base = smem_base + uint32_t(nstages) * stage_bytes;
  const uint32_t bar_base = scale_base + uint32_t(kTrMaxStages) * kTrScaleSlot;    // 16-byte aligned
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kTrMaxStages + s); };
  auto sfull_bar = [&](int s) { return bar_base + 8u * (2 * kTrMaxStages + s); };
  auto sempty_bar = [&](int s) { return bar_base + 8u * (3 * kTrMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (4 * kTrMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (4 * kTrMaxStages + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (4 * kTrMaxStages + 4);
  static_assert(kTrMaxStages * kTrScaleSlot + 8 * (4 * kTrMaxStages + 4) + 16 <= kTrTailBytes, "tail region");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_w);
    ptx::prefetch_tensormap(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < nstages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
      ptx::mbar_init(sfull_bar(s), 32);
      ptx::mbar_init(sempty_bar(s), 8);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(tfull_bar(s), 1);
      ptx::mbar_init(tempty_bar(s), 8);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc<1>(tmem_slot, 512);
    ptx::tmem_relinquish<1>();
  }
  if (warp == 3) {
    for (uint32_t off = uint32_t(lane) * 4u; off < uint32_t(nstages) * kTrScaleSlot; off += 128u)
      *reinterpret_cast<float*>(smem_raw + (scale_base - smem_base) + off) = 0.0f;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_base));

  const int num_kb = p.num_k_blocks;
  const int row0 = int(blockIdx.x) * p.rows_per_worker;
  const int row_end = min(row0 + p.rows_per_worker, int(p.N));

  if (warp == 0) {
    // =========================== TMA producer ===========================
    ptx::reg_dealloc<40>();
    const uint64_t pol_w = ptx::policy_evict_first();
    const uint64_t pol_x = ptx::policy_evict_last();
    int stage = 0;
    uint32_t phase = 0;
    for (int pass = 0; pass < p.passes; ++pass) {
      const int base = row0 + pass * p.rows_per_pass;
      if (base >= row_end) break;
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(empty_bar(stage), phase ^ 1);
        if (ptx::elect_one()) {
          const uint32_t dst = smem_base + uint32_t(stage) * stage_bytes;
          ptx::mbar_arrive_expect_tx(full_bar(stage), p.stage_tx_bytes);
          ptx::tma_load_2d(&tmap_w, full_bar(stage), dst, kb * BLOCK_K, base, pol_w);
          ptx::tma_load_2d(&tmap_x, full_bar(stage), dst + p.x_off, kb * BLOCK_K, 0, pol_x);
        }
        __syncwarp();
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    ptx::reg_dealloc<40>();
    const uint32_t tmem_base_u = __shfl_sync(0xFFFFFFFFu, tmem_base, 0);
    const uint64_t desc_hi = make_smem_desc(0);
    int stage = 0, slot = 0;
    uint32_t phase = 0, slot_phase = 0;
    for (int pass = 0; pass < p.passes; ++pass) {
      if (row0 + pass * p.rows_per_pass >= row_end) break;
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(full_bar(stage), phase);
        const uint32_t w_addr = smem_base + uint32_t(stage) * stage_bytes;
        const uint64_t db = desc_hi | uint64_t((w_addr & 0x3FFFFu) >> 4);                 // B operand: the weight rows
        const uint64_t da = desc_hi | uint64_t(((w_addr + p.x_off) & 0x3FFFFu) >> 4);     // A operand: the tokens
        ptx::mbar_wait(tempty_bar(slot), slot_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base_u + uint32_t(slot * 256);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            ptx::mma_f8f6f4<1>(d_tmem, da + uint64_t(k * (UMMA_K >> 4)), db + uint64_t(k * (UMMA_K >> 4)), p.idesc, k > 0 ? 1u : 0u);
          ptx::tc_commit<1>(empty_bar(stage));
          ptx::tc_commit<1>(tfull_bar(slot));
        }
        __syncwarp();
        if (++stage == nstages) { stage = 0; phase ^= 1; }
        if (++slot == 2) { slot = 0; slot_phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    ptx::reg_dealloc<40>();
  } else if (warp == 3) {
    // =========================== scale loader ===========================
    ptx::reg_dealloc<40>();
    int stage = 0;
    uint32_t phase = 0;
    for (int pass = 0; pass < p.passes; ++pass) {
      const int base = row0 + pass * p.rows_per_pass;
      if (base >= row_end) break;
      const int sbc = min(base / 128 + min(lane, 3), p.num_sb_cols - 1);
      const float* sb_ptr = p.sb + int64_t(sbc) * p.sb_sn;
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(sempty_bar(stage), phase ^ 1);
        const uint32_t slot = scale_base + uint32_t(stage) * kTrScaleSlot;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int t = lane + 32 * q;
          if (t < p.M) ptx::cp_async4(slot + 4u * t, p.sa + int64_t(t) * p.sa_sm + int64_t(kb) * p.sa_sk);
        }
        if (lane < 4) ptx::cp_async4(slot + 4u * 128 + 4u * lane, sb_ptr + int64_t(kb) * p.sb_sk);
        asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(sfull_bar(stage)) : "memory");
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
    }
    ptx::cp_async_commit();
    ptx::cp_async_wait<0>();
    __syncwarp();
  } else {
    // =========================== promotion + epilogue warps ===========================
    ptx::reg_alloc<232>();
    const int ew = warp - 4;
    const int quarter = warp & 3;                      // TMEM lanes this warp may touch = its 32 token rows
    const int cg = ew >> 2;                            // which 128-column half of the pass
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const int tok = quarter * 32 + lane;               // this thread's token row
    int stage = 0, slot = 0;
    uint32_t phase = 0, slot_phase = 0;
    OutT* Cp = reinterpret_cast<OutT*>(p.C);
    const bool vec_ok = (p.ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
    for (int pass = 0; pass < p.passes; ++pass) {
      const int base = row0 + pass * p.rows_per_pass;
      if (base >= row_end) break;
      const int limit = min(base + p.rows_per_pass, row_end);    // output columns [base, limit) belong to this pass
      const int n_first = base + cg * 128;                       // this warp's first output column
      const int cols_here = max(0, min(128, p.rows_per_pass - cg * 128));
      const int nch = (cols_here + 31) >> 5;                     // 32-column TMEM chunks this warp promotes (0..4)
      const int blk_off = n_first - (base / 128) * 128;          // column j lies in the pass's scale block (blk_off + j) / 128
      float accum[128];
#pragma unroll
      for (int i = 0; i < 128; ++i) accum[i] = 0.0f;
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(tfull_bar(slot), slot_phase);
        ptx::tc_fence_after();
        ptx::mbar_wait(sfull_bar(stage), phase);
        const uint32_t fslot = scale_base + uint32_t(stage) * kTrScaleSlot;
        const float sa_t = ptx::lds_f32(fslot + 4u * uint32_t(tok));
        const float f0 = sa_t * ptx::lds_f32(fslot + 512u), f1 = sa_t * ptx::lds_f32(fslot + 516u),
                    f2 = sa_t * ptx::lds_f32(fslot + 520u);
        const uint32_t taddr = tmem_base + lane_base + uint32_t(slot * 256 + cg * 128);
        auto release_slot = [&]() {                    // every tcgen05.ld of this warp on the slot has completed
          ptx::tc_fence_before();
          if (lane == 0) ptx::mbar_arrive(tempty_bar(slot));
        };
        uint32_t v[2][32];
        if (nch > 0) ptx::tmem_ld_32x32(taddr, v[0]);
        else release_slot();
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (c < nch) {                               // warp-uniform
            ptx::tmem_ld_wait(v[c & 1]);
            if (c + 1 < nch) ptx::tmem_ld_32x32(taddr + uint32_t((c + 1) * 32), v[(c + 1) & 1]);
            else release_slot();
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const int blk = (blk_off + c * 32 + g * 8) >> 7;
              const float s = blk == 0 ? f0 : (blk == 1 ? f1 : f2);
#pragma unroll
              for (int i = 0; i < 8; i += 2)
                ptx::fma2(accum[c * 32 + g * 8 + i], accum[c * 32 + g * 8 + i + 1], v[c & 1][g * 8 + i], v[c & 1][g * 8 + i + 1], s);
            }
          }
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(sempty_bar(stage));
        if (++stage == nstages) { stage = 0; phase ^= 1; }
        if (++slot == 2) { slot = 0; slot_phase ^= 1; }
      }
      // ---------------- this thread's token row x up to 128 consecutive columns ----------------
      if (tok < p.M) {
        OutT* crow = Cp + int64_t(tok) * p.ldc;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = n_first + j * 8;
          if (j * 8 < cols_here && col < limit) {
            const float* a = accum + j * 8;
            if (vec_ok && col + 8 <= limit) {
              uint4 w;
              w.x = pack2<OutT>(a[0], a[1]); w.y = pack2<OutT>(a[2], a[3]);
              w.z = pack2<OutT>(a[4], a[5]); w.w = pack2<OutT>(a[6], a[7]);
              *reinterpret_cast<uint4*>(crow + col) = w;
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i)
                if (col + i < limit) {
                  const uint32_t w = pack2<OutT>(a[i], 0.0f);
                  crow[col + i] = *reinterpret_cast<const OutT*>(&w);
                }
            }
          }
        }
      }
    }
    __syncwarp();
  }

  // =========================== teardown ===========================
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<1>(tmem_base, 512);
}

bool trows_supported(int64_t M, int64_t N, int64_t K) { return M >= 1 && M <= 128 && N >= 1 && K >= 1; }

template <typename OutT>
static int launch_trows(const CUtensorMap& tw, const CUtensorMap& tx, const TRowsParams& p, int workers, cudaStream_t st) {
  auto kern = fp8_blockwise_gemm_trows_kernel<OutT>;
  static std::atomic<uint64_t> configured{0};
  if (int e = ensure_max_dyn_smem(kern, configured, kMaxDynSmem)) return e;
  kern<<<dim3(unsigned(workers)), dim3(kTrThreads), kMaxDynSmem, st>>>(tw, tx, p);
  return int(cudaGetLastError());
}

// Arguments were validated by gemm_entry (pointers, strides, alignment, dtype); M in [1, 128].
int trows_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M, int64_t N, int64_t K,
                int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk, int64_t sb_sk, int64_t sb_sn,
                int out_dtype, int num_sms, cudaStream_t st) {
  if (num_sms <= 0) num_sms = device_sm_count();
  const int64_t W = num_sms;
  int64_t R = ((N + W - 1) / W + 7) / 8 * 8;                    // weight rows per CTA, whole 8-row swizzle atoms
  const int passes = int((R + 255) / 256);
  const int rows_per_pass = int(((R + passes - 1) / passes + 7) / 8 * 8);
  R = int64_t(rows_per_pass) * passes;
  const int workers = int((N + R - 1) / R);
  TRowsParams p;
  memset(&p, 0, sizeof(p));
  p.sa = sa; p.sb = sb; p.C = Cout;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.sa_sm = sa_sm; p.sa_sk = sa_sk; p.sb_sk = sb_sk; p.sb_sn = sb_sn;
  p.num_k_blocks = int32_t((K + BLOCK_K - 1) / BLOCK_K);
  p.num_sb_cols = int32_t((N + 127) / 128);
  p.rows_per_worker = int32_t(R);
  p.passes = passes;
  p.rows_per_pass = rows_per_pass;
  const uint32_t x_rows = uint32_t((M + 7) / 8 * 8);
  p.x_off = uint32_t(rows_per_pass) * BLOCK_K;                   // a multiple of 1 KB
  p.stage_bytes = p.x_off + BLOCK_M * BLOCK_K;                   // the MMA reads 128 token rows whatever the box held
  p.stage_tx_bytes = (uint32_t(rows_per_pass) + x_rows) * BLOCK_K;
  int n = (kMaxDynSmem - kTrTailBytes) / int(p.stage_bytes);
  p.num_stages = n > kTrMaxStages ? kTrMaxStages : n;
  if (p.num_stages < 2) return SGLK_E_SHAPE;
  p.idesc = make_idesc(BLOCK_M, (rows_per_pass + 15) / 16 * 16);
  CUtensorMap tw, tx;
  if (!make_tmap_2d(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(N), uint64_t(K), uint64_t(ldb), uint32_t(rows_per_pass), BLOCK_K))
    return SGLK_E_DRIVER;
  if (!make_tmap_2d(&tx, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, A, uint64_t(M), uint64_t(K), uint64_t(lda), x_rows, BLOCK_K))
    return SGLK_E_DRIVER;
  if (out_dtype == SGLK_BF16) return launch_trows<__nv_bfloat16>(tw, tx, p, workers, st);
  return launch_trows<__half>(tw, tx, p, workers, st);
}

}  // namespace sglk
