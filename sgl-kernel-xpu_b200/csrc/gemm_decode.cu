// Decode-shaped fp8 blockwise-scaled GEMMs for sm_100a (M <= 256 tokens): two "swap-AB" kernels.
//   1. fp8_blockwise_gemm_decode_kernel (first half of this file): 128 / 256-row weight TILES, stream-K hybrid schedule with a
//      workspace fix-up — variants swapab_cta1 / swapab_cta2, the default of early round 2, kept as explicit variants;
//   2. fp8_blockwise_gemm_rows_kernel (second half): one contiguous weight-ROW RANGE per CTA for all of K (or per cluster
//      for a K range), no workspace, no atomics — variant swapab_rows, the default up to 64 tokens (gemm.cu pick_variant).
//
//   C[m, n] = sum_kb sa[m, kb] * sb[kb, n / 128] * sum_{k in kb} A[m, k] * B[k, n]
//
// Same op as gemm.cu (declared /root/reference/include/sgl_kernel_ops.h:581-586, semantics
// tests/test_fp8_blockwise_gemm.py:23-63; the reference ships no kernel), for the shapes where it is a weight
// stream: N*K bytes of e4m3 weights pass through once, everything else is small.  Design:
//   * the WEIGHT rows are the MMA M side (128 rows per CTA, 256 per CTA pair) and the tokens the MMA N side
//     (NT = 16 / 32 / 64 / 128 / 256 columns): an accumulator stage is 128 lanes x NT columns of TMEM, so up to eight
//     stages fit and the MMA issuer never waits for the promotion warps; a K block's promotion is 128*NT FMAs instead
//     of 128*128; the weight scale sb[kb, n/128] is one scalar per (tile, K block); and a split-K partial is
//     128*NT*4 bytes (8 KB at 16 tokens) — so cutting tiles across CTAs (stream-K) to balance 148 SMs is nearly free.
//   * CTA-pair form (cta_group::2): the pair's MMA covers 256 weight rows x NT tokens, each CTA TMA-loads its own
//     128 weight rows and HALF of the activation tile — the L2 -> smem traffic of the activations (which every tile
//     re-reads) is halved, which is what limits single-CTA tiles from ~64 tokens up.
//   * per K block the activation scales sa[0..M, kb] and the tile's sb[kb] travel global -> smem with cp.async
//     (any strides), tracked by an mbarrier, so the promotion warps never wait on a global load.
//   * stream-K: the (tile, K block) units are cut into one contiguous range per CTA (pair); tiles shared by several
//     workers are summed through the workspace by whichever arrives last, in worker order (bit-reproducible).
//   * C^T leaves the registers directly: lane = output column, so two neighbouring lanes swap one value and each
//     writes 4 bytes; a warp store covers two 64-byte row segments.
#include "gemm_common.cuh"

namespace sglk {

// warp 0 TMA producer, 1 MMA issuer, 2 TMEM owner, 3 scale loader, 4.. promotion/epilogue warps: 8 (two per TMEM lane
// quarter, NT/2 token columns each) up to 32 tokens, 16 (NT/4 columns each) above — the promotion of a K block is a chain
// of barrier probe -> tcgen05.ld -> smem scale reads -> FMAs whose latencies only overlap across warps
constexpr int kDecMaxStages = 16;
constexpr int kWTileBytes = BLOCK_M * BLOCK_K;   // 16 KB of weights per CTA per K block

struct DecodeParams {
  const float* sa;
  const float* sb;
  void* C;
  int64_t M, N, K, ldc;
  int64_t sa_sm, sa_sk, sb_sk, sb_sn;
  int32_t num_tiles, num_k_blocks, num_sb_cols;
  int32_t stream_k;
  int32_t sk_tile_base;      // tiles [sk_tile_base, num_tiles) are cut across the workers (processed FIRST); tiles below it are
                             // whole tiles, worker w takes w, w + workers, ... (processed after); == num_tiles: no stream-K
  int32_t sk_workers;        // workers 0 .. sk_workers-1 share the stream-K units (every one of them gets at least one)
  int32_t num_stages;
  uint32_t stage_bytes;      // weight tile + activation slot (multiple of 1024)
  uint32_t scale_slot_bytes; // per-stage scale slot: NT floats of sa + sb (padded to 16 bytes)
  uint32_t x_rows[2];        // activation rows CTA rank r loads per K block (multiple of 8; 0 = none exist)
  uint32_t stage_tx_bytes;   // bytes the (leader's) full barrier sees per stage
  int32_t* ws_counters;      // [num_tiles][kCG * epilogue warps], zero on entry, left zero
  float* ws_partials;        // [2 * workers][kCG * epilogue warps][CW * 32]
};

namespace dptx {
template <int N>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&v)[N]);
template <>
__device__ __forceinline__ void tmem_ld<4>(uint32_t taddr, uint32_t (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr)
               : "memory");
}
template <>
__device__ __forceinline__ void tmem_ld<8>(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
template <>
__device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
template <>
__device__ __forceinline__ void tmem_ld<32>(uint32_t taddr, uint32_t (&v)[32]) {
  ptx::tmem_ld_32x32(taddr, v);
}
// wait for this thread's outstanding tcgen05.ld; the registers are threaded through so their uses stay below
template <int N>
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[N]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < N; ++i) asm volatile("" : "+r"(v[i]));
}
// {a0, a1} += {v0, v1} * {s0, s1}
__device__ __forceinline__ void fma2_vec(float& a0, float& a1, uint32_t v0, uint32_t v1, float s0, float s1) {
  asm("{\n\t.reg .b64 acc, val, scl;\n\t"
      "mov.b64 acc, {%0, %1};\n\t"
      "mov.b64 val, {%2, %3};\n\t"
      "mov.b64 scl, {%4, %5};\n\t"
      "fma.rn.f32x2 acc, val, scl, acc;\n\t"
      "mov.b64 {%0, %1}, acc;\n\t}"
      : "+f"(a0), "+f"(a1)
      : "r"(v0), "r"(v1), "f"(s0), "f"(s1));
}
// {a0, a1} += {v0, v1} * ({s0, s1} * {t, t})   — two packed fp32 ops
__device__ __forceinline__ void fma2_scaled(float& a0, float& a1, uint32_t v0, uint32_t v1, float s0, float s1, float t) {
  asm("{\n\t.reg .b64 acc, val, scl, tt;\n\t"
      "mov.b64 acc, {%0, %1};\n\t"
      "mov.b64 val, {%2, %3};\n\t"
      "mov.b64 scl, {%4, %5};\n\t"
      "mov.b64 tt, {%6, %6};\n\t"
      "mul.rn.f32x2 scl, scl, tt;\n\t"
      "fma.rn.f32x2 acc, val, scl, acc;\n\t"
      "mov.b64 {%0, %1}, acc;\n\t}"
      : "+f"(a0), "+f"(a1)
      : "r"(v0), "r"(v1), "f"(s0), "f"(s1), "f"(t));
}
// every prior cp.async of this thread has landed -> one (pre-counted) arrival on the mbarrier
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
// tcgen05.mma (cta_group::1, e4m3 x e4m3 -> f32) with the operand descriptors given as 32-bit halves: the low words
// (start address >> 4; leading byte offset 0) differ per tile, the high word (stride offset, version, swizzle) is shared
__device__ __forceinline__ void mma_f8_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], da, db, %4, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ float4 lds_f32x4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
}  // namespace dptx

// One worker's walk over its share of the problem.  Hybrid schedule: the tiles that do not fill a whole wave
// ([sk_tile_base, num_tiles)) are cut into one contiguous range of (tile, K block) units per worker and come FIRST — their
// partials are published, and the last arriver's fix-up runs, while the TMA / MMA pipeline is already streaming the
// worker's whole tiles, so no fix-up latency is left at the end of the kernel; then whole tiles w, w + workers, ...
struct DecIter {
  int64_t cur, end;
  int num_kb, tile_base, next_whole, step;
  __device__ DecIter(const DecodeParams& p, int worker, int num_workers) {
    num_kb = p.num_k_blocks;
    tile_base = p.sk_tile_base;
    const int64_t total = int64_t(p.num_tiles - tile_base) * num_kb;
    cur = worker < p.sk_workers ? total * worker / p.sk_workers : 0;
    end = worker < p.sk_workers ? total * (worker + 1) / p.sk_workers : 0;
    next_whole = worker;
    step = num_workers;
  }
  __device__ bool next(int& tile, int& kb0, int& kb1) {
    if (cur < end) {
      const int t = int(cur / num_kb);
      tile = tile_base + t;
      kb0 = int(cur - int64_t(t) * num_kb);
      const int64_t left = end - cur;
      kb1 = left < int64_t(num_kb - kb0) ? kb0 + int(left) : num_kb;
      cur += kb1 - kb0;
      return true;
    }
    if (next_whole < tile_base) {
      tile = next_whole; kb0 = 0; kb1 = num_kb;
      next_whole += step;
      return true;
    }
    return false;
  }
};

template <int NT, int kCG>
struct DecCfg {
  static constexpr int kEW = 8;                              // promotion / epilogue warps: two per TMEM lane quarter
  static constexpr int kThreads = 128 + kEW * 32;
  static constexpr int CW = NT / 2;                          // token columns per promotion warp
  // Accumulator hand-offs are kHB K blocks deep: the MMA issuer fills kHB column ranges of one TMEM slot and commits
  // ONCE, the promotion warps wait once, pull the blocks out and fold them in.  The promotion of a hand-off is a chain of
  // latencies (barrier probe, tcgen05.ld, smem factor reads, FMAs) that a warp walks in order: per K block it measured
  // ~700-1200 clk against the ~690 clk at which a block's 16 KB of weights arrive, i.e. the kernel was bound by that
  // chain, not by HBM; per hand-off of 2-4 blocks it is amortised.
  static constexpr int kHB = NT <= 32 ? 4 : (NT <= 128 ? 2 : 1);
  static constexpr int kSlots = (512 / (NT * kHB)) < 8 ? (512 / (NT * kHB)) : 8;   // hand-off slots in TMEM
  static constexpr int TMEM_COLS = kSlots * kHB * NT;
  static constexpr int IW = CW < 32 ? CW : 32;               // columns per tcgen05.ld item
  static constexpr int IPB = CW / IW;                        // items per K block (per warp)
  static constexpr int ITEMS = kHB * IPB;                    // items per hand-off
  static constexpr bool kOneBatch = NT <= 64;                // the whole hand-off is pulled with one tcgen05.wait::ld
  static constexpr int NBUF = kOneBatch ? ITEMS : 2;
  static constexpr int X_ROWS = NT / kCG;                    // activation rows in one CTA's smem slot
  static constexpr int X_SLOT = (X_ROWS * BLOCK_K + 1023) / 1024 * 1024;
  static constexpr int STAGE_BYTES = kWTileBytes + X_SLOT;
  static constexpr int SCALE_SLOT = NT * 4 + 16;
  // setmaxnreg budgets: the CTA owns kThreads x 168 registers from launch; the sum over the warps after the re-split
  // must not exceed that, or setmaxnreg.inc waits forever
  static constexpr int kCtlRegs = 40;
  static constexpr int kEpiRegs = 232;
  static_assert(4 * 32 * kCtlRegs + kEW * 32 * kEpiRegs <= kThreads * 168, "register re-split");
  static_assert(NT % 16 == 0 && NT >= 16 && NT <= 256, "MMA N");
  static_assert((TMEM_COLS & (TMEM_COLS - 1)) == 0 && TMEM_COLS >= 32 && TMEM_COLS <= 512, "TMEM columns");
};

template <int NT, int kCG, typename OutT>
__global__ void __launch_bounds__(DecCfg<NT, kCG>::kThreads, 1)
fp8_blockwise_gemm_decode_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x0,
                                 const __grid_constant__ CUtensorMap tmap_x1, const DecodeParams p) {
  using C = DecCfg<NT, kCG>;
  constexpr int CW = C::CW, kHB = C::kHB, kSlots = C::kSlots, kEW = C::kEW;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  if ((ptx::smem_u32(smem_raw) & 1023u) != 0) __trap();      // 128B-swizzled tiles need 1 KB alignment; no slack is budgeted
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  const int nstages = p.num_stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t scale_base = smem_base + uint32_t(nstages) * stage_bytes;
  const uint32_t bar_base = scale_base + uint32_t(nstages) * p.scale_slot_bytes;   // 16-byte aligned
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kDecMaxStages + s); };
  auto sfull_bar = [&](int s) { return bar_base + 8u * (2 * kDecMaxStages + s); };
  auto sempty_bar = [&](int s) { return bar_base + 8u * (3 * kDecMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (4 * kDecMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (4 * kDecMaxStages + 8 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (4 * kDecMaxStages + 16);
  uint8_t* smem_gen = smem_raw;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (kCG == 2) ? ptx::cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_w);
    ptx::prefetch_tensormap(&tmap_x0);
    if (kCG == 2) ptx::prefetch_tensormap(&tmap_x1);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < nstages; ++s) {
      ptx::mbar_init(full_bar(s), 1);                  // the (leader's) producer's arrive.expect_tx
      ptx::mbar_init(empty_bar(s), 1);                 // one tcgen05.commit
      ptx::mbar_init(sfull_bar(s), 32);                // every lane of the scale loader: cp.async.mbarrier.arrive.noinc
      ptx::mbar_init(sempty_bar(s), kEW);              // this CTA's promotion warps have read the slot
    }
    for (int s = 0; s < kSlots; ++s) {
      ptx::mbar_init(tfull_bar(s), 1);                 // one tcgen05.commit per hand-off
      ptx::mbar_init(tempty_bar(s), kCG * kEW);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc<kCG>(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish<kCG>();
  }
  if (warp == 3) {
    // factor slots of tokens >= M are never filled: give them a defined value (their columns are never stored)
    for (uint32_t off = uint32_t(lane) * 4u; off < uint32_t(nstages) * p.scale_slot_bytes; off += 128u)
      *reinterpret_cast<float*>(smem_gen + (scale_base - smem_base) + off) = 0.0f;
  }
  ptx::tc_fence_before();
  if (kCG == 2) ptx::cluster_sync(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  const int num_workers = gridDim.x / kCG;
  const int worker = blockIdx.x / kCG;
  const int num_kb = p.num_k_blocks;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    ptx::reg_dealloc<C::kCtlRegs>();
    const uint64_t pol_w = ptx::policy_evict_first();      // weights pass through once
    const uint64_t pol_x = ptx::policy_evict_last();       // the activation K blocks are re-read by every tile
    const uint32_t x_rows = p.x_rows[cta_rank];
    const CUtensorMap* tmap_x = (kCG == 2 && cta_rank == 1) ? &tmap_x1 : &tmap_x0;
    int stage = 0;
    uint32_t phase = 0;
    DecIter it(p, worker, num_workers);
    int tile, kb0, kb1;
    while (it.next(tile, kb0, kb1)) {
      const int n0 = tile * (BLOCK_M * kCG) + int(cta_rank) * BLOCK_M;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(empty_bar(stage), phase ^ 1);
        const uint32_t w_addr = smem_base + stage * stage_bytes;
        const uint32_t x_addr = w_addr + kWTileBytes;
        if (ptx::elect_one()) {
          if constexpr (kCG == 1) {
            ptx::mbar_arrive_expect_tx(full_bar(stage), p.stage_tx_bytes);
            ptx::tma_load_2d(&tmap_w, full_bar(stage), w_addr, kb * BLOCK_K, n0, pol_w);
            ptx::tma_load_2d(tmap_x, full_bar(stage), x_addr, kb * BLOCK_K, 0, pol_x);
          } else {
            const uint32_t leader_full = full_bar(stage) & 0xFEFFFFFFu;   // both CTAs' bytes are counted on the leader's barrier
            if (is_leader) ptx::mbar_arrive_expect_tx(full_bar(stage), p.stage_tx_bytes);
            ptx::tma_load_2d_pair(&tmap_w, leader_full, w_addr, kb * BLOCK_K, n0, pol_w);
            if (x_rows) ptx::tma_load_2d_pair(tmap_x, leader_full, x_addr, kb * BLOCK_K, int(cta_rank) * C::X_ROWS, pol_x);
          }
        }
        __syncwarp();
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer (leader CTA only) ===========================
    ptx::reg_dealloc<C::kCtlRegs>();
    if (is_leader) {
      constexpr uint32_t idesc = make_idesc(BLOCK_M * kCG, NT);
      const uint32_t tmem_base_u = __shfl_sync(0xFFFFFFFFu, tmem_base, 0);
      const uint64_t desc_hi = make_smem_desc(0);
      int stage = 0, slot = 0;
      uint32_t phase = 0, slot_phase = 0;
      DecIter it(p, worker, num_workers);
      int tile, kb0, kb1;
      while (it.next(tile, kb0, kb1)) {
        int j = 0;                                     // position of this K block inside its hand-off
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(full_bar(stage), phase);
          const uint32_t w_addr = smem_base + stage * stage_bytes;
          const uint64_t da = desc_hi | uint64_t((w_addr & 0x3FFFFu) >> 4);
          const uint64_t db = desc_hi | uint64_t(((w_addr + kWTileBytes) & 0x3FFFFu) >> 4);
          if (j == 0) ptx::mbar_wait(tempty_bar(slot), slot_phase ^ 1);   // slot drained by every promotion warp (of both CTAs)?
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base_u + uint32_t((slot * kHB + j) * NT);
          const bool last = (j == kHB - 1) || (kb + 1 == kb1);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
              ptx::mma_f8f6f4<kCG>(d_tmem, da + uint64_t(k * (UMMA_K >> 4)), db + uint64_t(k * (UMMA_K >> 4)), idesc,
                                   k > 0 ? 1u : 0u);
            ptx::tc_commit<kCG>(empty_bar(stage));     // smem stage reusable once these MMAs have read it
            if (last) ptx::tc_commit<kCG>(tfull_bar(slot));   // the hand-off's partials are complete
          }
          __syncwarp();
          if (++stage == nstages) { stage = 0; phase ^= 1; }
          if (last) {
            j = 0;
            if (++slot == kSlots) { slot = 0; slot_phase ^= 1; }
          } else {
            ++j;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    ptx::reg_dealloc<C::kCtlRegs>();                   // TMEM owner: only waits for the end
  } else if (warp == 3) {
    // =========================== scale loader ===========================
    // sa[token, kb] for every token and this CTA's sb[kb, n0 / 128] -> the stage's factor slot, 4 bytes per cp.async (any
    // strides), completion tracked by the slot's mbarrier.  Fully asynchronous: as many K blocks in flight as the ring
    // has stages.  (A register-staged version — load, multiply by sb, st.shared — could only run one L2 round trip
    // ahead and capped every CTA at one K block per ~0.2-0.5 us: slower than the weights arrive.)
    ptx::reg_dealloc<C::kCtlRegs>();
    int stage = 0;
    uint32_t phase = 0;
    DecIter it(p, worker, num_workers);
    int tile, kb0, kb1;
    while (it.next(tile, kb0, kb1)) {
      const int n0 = tile * (BLOCK_M * kCG) + int(cta_rank) * BLOCK_M;
      const int sbc = min(n0 / 128, p.num_sb_cols - 1);
      const float* sb_ptr = p.sb + int64_t(sbc) * p.sb_sn;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(sempty_bar(stage), phase ^ 1);
        const uint32_t slot = scale_base + uint32_t(stage) * p.scale_slot_bytes;
#pragma unroll
        for (int j = 0; j < (NT + 31) / 32; ++j) {
          const int t = lane + 32 * j;
          if (t < NT && t < p.M) ptx::cp_async4(slot + 4u * t, p.sa + int64_t(t) * p.sa_sm + int64_t(kb) * p.sa_sk);
        }
        if (lane == 0) ptx::cp_async4(slot + 4u * NT, sb_ptr + int64_t(kb) * p.sb_sk);
        dptx::cp_async_mbar_arrive_noinc(sfull_bar(stage));
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
    }
    ptx::cp_async_commit();
    ptx::cp_async_wait<0>();
    __syncwarp();
  } else {
    // =========================== promotion + epilogue warps ===========================
    ptx::reg_alloc<C::kEpiRegs>();
    const int ew = warp - 4;
    const int quarter = warp & 3;                      // TMEM lanes this warp may touch: 32*quarter .. +31
    const int colgroup = ew >> 2;                      // which half of the token columns
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const uint32_t tempty_leader = (kCG == 2) ? ptx::mapa(tempty_bar(0), 0) : tempty_bar(0);
    int stage = 0, slot = 0;
    uint32_t phase = 0, slot_phase = 0;
    DecIter it(p, worker, num_workers);
    int tile, kb0, kb1;
    bool first_segment = true;
    const int t0 = colgroup * CW;                      // first token column of this warp
    constexpr int IW = C::IW, IPB = C::IPB, ITEMS = C::ITEMS, NBUF = C::NBUF;
    while (it.next(tile, kb0, kb1)) {
      const int n = tile * (BLOCK_M * kCG) + int(cta_rank) * BLOCK_M + quarter * 32 + lane;   // this lane's output column
      float accum[CW];
#pragma unroll
      for (int i = 0; i < CW; ++i) accum[i] = 0.0f;
      for (int kb = kb0; kb < kb1; kb += kHB) {
        const int nb = min(kHB, kb1 - kb);             // K blocks in this hand-off (the last one of a segment may be short)
        // ring position of each of its blocks
        int st[kHB];
        uint32_t ph[kHB];
#pragma unroll
        for (int j = 0; j < kHB; ++j) {
          int sj = stage + j;
          const bool wrap = sj >= nstages;
          st[j] = wrap ? sj - nstages : sj;
          ph[j] = wrap ? phase ^ 1 : phase;
        }
        ptx::mbar_wait(tfull_bar(slot), slot_phase);   // the hand-off's MMAs complete
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + lane_base + uint32_t(slot * kHB * NT + t0);
        uint32_t v[NBUF][IW];
        auto item_ld = [&](int i, uint32_t (&buf)[IW]) {      // item i = (block i / IPB, column chunk i % IPB)
          dptx::tmem_ld<IW>(taddr + uint32_t((i / IPB) * NT + (i % IPB) * IW), buf);
        };
        auto item_fma = [&](int i, const uint32_t (&buf)[IW]) {
          const uint32_t slot_addr = scale_base + uint32_t(st[i / IPB]) * p.scale_slot_bytes;
          const uint32_t fac = slot_addr + 4u * uint32_t(t0 + (i % IPB) * IW);
          const float sbv = ptx::lds_f32(slot_addr + 4u * NT);
#pragma unroll
          for (int q = 0; q < IW / 4; ++q) {
            const float4 f = dptx::lds_f32x4(fac + 16u * q);
            float* a = accum + (i % IPB) * IW + 4 * q;
            dptx::fma2_scaled(a[0], a[1], buf[4 * q], buf[4 * q + 1], f.x, f.y, sbv);
            dptx::fma2_scaled(a[2], a[3], buf[4 * q + 2], buf[4 * q + 3], f.z, f.w, sbv);
          }
          if ((i % IPB) == IPB - 1) {                  // every factor of this block has been read: its slot can be refilled
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(sempty_bar(st[i / IPB]));
          }
        };
        auto release_slot = [&]() {                    // every tcgen05.ld of this warp on the slot has completed
          ptx::tc_fence_before();
          if (lane == 0) {
            if constexpr (kCG == 2) ptx::mbar_arrive_cluster(tempty_leader + 8u * slot);
            else ptx::mbar_arrive(tempty_bar(slot));
          }
        };
        if constexpr (C::kOneBatch) {
          // the whole hand-off in one go: all tcgen05.ld issued, one wait, slot handed back, then the FMAs
#pragma unroll
          for (int i = 0; i < ITEMS; ++i)
            if (i / IPB < nb) item_ld(i, v[i]);
#pragma unroll
          for (int j = 0; j < kHB; ++j)
            if (j < nb) ptx::mbar_wait(sfull_bar(st[j]), ph[j]);     // factors of block j in smem (long since, normally)
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          release_slot();
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) {
            if (i / IPB < nb) {
#pragma unroll
              for (int r = 0; r < IW; ++r) asm volatile("" : "+r"(v[i][r]));   // consumers stay below the wait
              item_fma(i, v[i]);
            }
          }
        } else {
          // item i+1 is in flight while item i is folded in
          ptx::mbar_wait(sfull_bar(st[0]), ph[0]);
          item_ld(0, v[0]);
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) {
            if (i / IPB < nb) {
              dptx::tmem_ld_wait<IW>(v[i & 1]);
              const bool more = (i + 1 < ITEMS) && ((i + 1) / IPB < nb);
              if (more) {
                if ((i + 1) % IPB == 0) ptx::mbar_wait(sfull_bar(st[(i + 1) / IPB]), ph[(i + 1) / IPB]);
                item_ld(i + 1, v[(i + 1) & 1]);
              } else {
                release_slot();
              }
              item_fma(i, v[i & 1]);
            }
          }
        }
        stage += nb;
        if (stage >= nstages) { stage -= nstages; phase ^= 1; }
        if (++slot == kSlots) { slot = 0; slot_phase ^= 1; }
      }
      // ---------------- stream-K: a tile shared with other workers is completed by the last arriver ----------------
      bool write_out = true;
      if (p.stream_k && tile >= p.sk_tile_base && (kb0 != 0 || kb1 != num_kb)) {
        const int tile_rel = tile - p.sk_tile_base;
        const int64_t total = int64_t(p.num_tiles - p.sk_tile_base) * num_kb;
        const int skw = p.sk_workers;
        auto range_begin = [&](int w) { return total * w / skw; };
        auto owner = [&](int64_t u) { return int(((u + 1) * skw + total - 1) / total) - 1; };
        const int c_first = owner(int64_t(tile_rel) * num_kb), c_last = owner(int64_t(tile_rel + 1) * num_kb - 1);
        auto slot_of = [&](int c) { return 2 * c + ((int(range_begin(c) / num_kb) == tile_rel) ? 0 : 1); };
        constexpr int kWarpFloats = CW * 32;
        auto part_ptr = [&](int c) {
          return p.ws_partials + (int64_t(slot_of(c)) * (kCG * kEW) + int(cta_rank) * kEW + ew) * kWarpFloats;
        };
        float4* mine = reinterpret_cast<float4*>(part_ptr(worker));
        int32_t* cnt = p.ws_counters + int64_t(tile) * (kCG * kEW) + int(cta_rank) * kEW + ew;
        const int n_others = c_last - c_first;
        // the worker that holds a tile's FIRST K blocks reaches it after the others started on theirs: it peeks at the
        // counter and, if everybody is in, skips publishing its own partial (three L2 round trips) and folds the others
        // into its registers in worker order
        bool peeked = false;
        if (!first_segment && kb0 == 0) {
          int seen = 0;
          if (lane == 0) seen = *reinterpret_cast<volatile int32_t*>(cnt);
          peeked = __shfl_sync(0xFFFFFFFFu, seen, 0) == n_others;
        }
        bool finisher = peeked;
        if (!peeked) {
#pragma unroll
          for (int j = 0; j < CW / 4; ++j)
            __stcg(mine + j * 32 + lane, make_float4(accum[4 * j], accum[4 * j + 1], accum[4 * j + 2], accum[4 * j + 3]));
          __syncwarp();
          int arrived = 0;
          if (lane == 0) {
            asm volatile("fence.acq_rel.gpu;" ::: "memory");   // (not __threadfence(): that is the SC fence, ~6 us here)
            arrived = atomicAdd(cnt, 1);
          }
          arrived = __shfl_sync(0xFFFFFFFFu, arrived, 0);
          finisher = arrived == n_others;                 // everyone else's partial is already visible
        }
        write_out = finisher;
        if (write_out) {
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          if (lane == 0) *cnt = 0;                        // leave the counters zeroed for the next launch
          // fixed order (bit-reproducible): 0 + partial(c_first) + partial(c_first + 1) + ...; the loads of a batch of
          // contributors are issued together, the additions stay in worker order
#pragma unroll
          for (int i = 0; i < CW; ++i) accum[i] = peeked ? 0.0f + accum[i] : 0.0f;
          constexpr int kBatch = CW <= 8 ? 4 : (CW <= 16 ? 2 : 1);
          for (int c = peeked ? c_first + 1 : c_first; c <= c_last; c += kBatch) {
            float4 t[kBatch][CW / 4];
#pragma unroll
            for (int b = 0; b < kBatch; ++b) {
              if (c + b <= c_last) {
                const float4* src = reinterpret_cast<const float4*>(part_ptr(c + b));
#pragma unroll
                for (int j = 0; j < CW / 4; ++j) t[b][j] = __ldcg(src + j * 32 + lane);
              }
            }
#pragma unroll
            for (int b = 0; b < kBatch; ++b) {
              if (c + b <= c_last) {
#pragma unroll
                for (int j = 0; j < CW / 4; ++j) {
                  accum[4 * j] += t[b][j].x; accum[4 * j + 1] += t[b][j].y;
                  accum[4 * j + 2] += t[b][j].z; accum[4 * j + 3] += t[b][j].w;
                }
              }
            }
          }
        }
      }
      first_segment = false;
      if (!write_out) continue;
      // ---------------- C^T out of the registers: lane = output column n, accum[j] = token t0 + j ----------------
      OutT* Cp = reinterpret_cast<OutT*>(p.C);
      const bool pair_ok = (p.ldc % 2 == 0) && (p.N % 2 == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 3) == 0);
      if (pair_ok) {
        // neighbouring lanes swap one value each: the even lane ends up with (n, n+1) of token j, the odd lane with
        // (n-1, n) of token j+1 — one 4-byte store per two values, 64-byte row segments per warp store
        const int odd = lane & 1;
        const int col = n - odd;
#pragma unroll
        for (int j = 0; j < CW; j += 2) {
          const float send = odd ? accum[j] : accum[j + 1];
          const float recv = __shfl_xor_sync(0xFFFFFFFFu, send, 1);
          const uint32_t w = odd ? pack2<OutT>(recv, accum[j + 1]) : pack2<OutT>(accum[j], recv);
          const int tok = t0 + j + odd;
          if (tok < p.M && col < p.N)
            *reinterpret_cast<uint32_t*>(Cp + int64_t(tok) * p.ldc + col) = w;
        }
      } else if (n < p.N) {
#pragma unroll
        for (int j = 0; j < CW; ++j) {
          if (t0 + j < p.M) {
            const uint32_t w = pack2<OutT>(accum[j], 0.0f);
            Cp[int64_t(t0 + j) * p.ldc + n] = *reinterpret_cast<const OutT*>(&w);
          }
        }
      }
    }
    __syncwarp();
  }

  // =========================== teardown ===========================
  ptx::tc_fence_before();
  if (kCG == 2) ptx::cluster_sync(); else __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<kCG>(tmem_base, C::TMEM_COLS);
}

// --------------------------------------------------------------------------------------------
// host side
// --------------------------------------------------------------------------------------------
static int decode_nt(int64_t M) { return M <= 16 ? 16 : (M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256))); }

bool decode_supported(int64_t M, int64_t N, int64_t K) { return M >= 1 && M <= 256 && N >= 1 && K >= 1; }

struct DecodePlan {
  int nt, cg, workers, sk_workers, stream_k, num_tiles, num_kb, stages, sk_tile_base;
  uint32_t stage_bytes, scale_slot;
  int64_t partial_bytes;
};

// Schedules, in K-block times per worker (G = K blocks per tile, W workers, T tiles):
//   whole tiles only               : ceil(T / W) * G
//   hybrid (stream-K remainder)    : floor(T / W) * G + ceil((T mod W) * G / W) + fix-up
// The fix-up (a partial of 128 x NT fp32 per CTA published, the last arriver re-reads the others') runs FIRST and
// overlaps the whole tiles that follow, so it is cheap when there is a whole tile behind it (N = 24576: one per worker)
// and costs its full latency — ~2 us ~ 6 K-block times — when there is none (N = 2112: 17 tiles for 148 CTAs).
// flags bits 16..23 (tuning hook): workers per tile when T < W, 0 = as many as there are.
static DecodePlan make_decode_plan(int64_t M, int64_t N, int64_t K, int cg, int num_sms, bool have_ws, int flags) {
  DecodePlan pl;
  pl.nt = decode_nt(M);
  pl.cg = cg;
  pl.num_tiles = int((N + BLOCK_M * cg - 1) / (BLOCK_M * cg));
  pl.num_kb = int((K + BLOCK_K - 1) / BLOCK_K);
  int w0 = num_sms / cg;
  if (w0 < 1) w0 = 1;
  const int kew = 8;
  const int T = pl.num_tiles, G = pl.num_kb;
  const int whole = T / w0, rem = T % w0;
  const int64_t rem_units = int64_t(rem) * G;
  // Schedule, from measurements in a CUDA graph on B200 (tools/decode_bench.py, tools/decode_split.py; K=7168):
  //  * at least one tile per worker (T >= W, e.g. N=24576): whole tiles.  The kernel is HBM-bound there and a second,
  //    partly filled wave still saturates the memory system (pair kernel, 96 tiles on 74 pairs: 35.8 us whole against
  //    39.5 us with the remainder tiles cut stream-K style — the fix-up traffic costs more than the balance gains);
  //  * fewer tiles than workers (N <= 9472 / 18944): each tile is split over min(W / T, 4) workers — more than four
  //    partial tiles per finisher cost more than the shorter K ranges gain (N=2112, M=16: 15.0 / 12.8 / 11.3 / 10.8 /
  //    10.9 / 10.9 / 14.0 us at 1 / 2 / 3 / 4 / 6 / 8 / 12 workers per tile).
  // SGLK_GEMM_STREAMK_FORCE keeps the general hybrid (remainder tiles split over all workers) for tests and tuning,
  // and bits 16..23 of flags override the workers per tile.
  const int sk_mode = (flags & SGLK_GEMM_STREAMK_FORCE) ? 1 : ((flags & SGLK_GEMM_STREAMK_NEVER) ? -1 : 0);
  int sk_workers = w0;
  bool want = sk_mode > 0;
  if (whole == 0) {
    int per_tile = (flags >> 16) & 0xFF;
    if (per_tile == 0 && sk_mode == 0) per_tile = w0 / T > 4 ? 4 : w0 / T;
    if (per_tile > 0 && int64_t(T) * per_tile < sk_workers) sk_workers = T * per_tile;
    if (sk_mode == 0) want = sk_workers > T;
  }
  if (rem_units < sk_workers) sk_workers = int(rem_units);   // every stream-K worker owns at least one unit
  const bool possible = have_ws && G >= 2 && rem > 0 && int64_t(T) * cg * kew * 4 <= kCounterAreaBytes;
  pl.stream_k = (possible && sk_mode >= 0 && want) ? 1 : 0;
  pl.sk_workers = pl.stream_k ? sk_workers : 0;
  pl.workers = pl.stream_k ? (whole > 0 ? w0 : sk_workers) : (T < w0 ? T : w0);
  if (pl.workers < 1) pl.workers = 1;
  pl.sk_tile_base = pl.stream_k ? whole * pl.workers : T;
  pl.partial_bytes = int64_t(2) * pl.workers * cg * 128 * pl.nt * 4;
  const int x_slot = ((pl.nt / cg) * BLOCK_K + 1023) / 1024 * 1024;
  pl.stage_bytes = uint32_t(kWTileBytes + x_slot);
  pl.scale_slot = uint32_t(pl.nt * 4 + 16);
  const int bar_bytes = 8 * (4 * kDecMaxStages + 16) + 16;
  int n = (kMaxDynSmem - bar_bytes) / int(pl.stage_bytes + pl.scale_slot);
  pl.stages = n > kDecMaxStages ? kDecMaxStages : n;
  return pl;
}

int64_t decode_workspace_bytes(int64_t M, int64_t N, int64_t K, int cg, int num_sms, int flags) {
  const DecodePlan pl = make_decode_plan(M, N, K, cg, num_sms, true, flags);
  return pl.stream_k ? kCounterAreaBytes + pl.partial_bytes : 0;
}

template <int NT, int kCG, typename OutT>
static int launch_decode(const CUtensorMap& tw, const CUtensorMap& tx0, const CUtensorMap& tx1, const DecodeParams& p,
                         int workers, cudaStream_t st) {
  auto kern = fp8_blockwise_gemm_decode_kernel<NT, kCG, OutT>;
  static std::atomic<uint64_t> configured{0};
  if (int e = ensure_max_dyn_smem(kern, configured, kMaxDynSmem)) return e;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(unsigned(workers * kCG));
  cfg.blockDim = dim3(DecCfg<NT, kCG>::kThreads);
  cfg.dynamicSmemBytes = kMaxDynSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return int(cudaLaunchKernelEx(&cfg, kern, tw, tx0, tx1, p));
}

template <int kCG, typename OutT>
static int launch_decode_nt(int nt, const CUtensorMap& tw, const CUtensorMap& tx0, const CUtensorMap& tx1,
                            const DecodeParams& p, int workers, cudaStream_t st) {
  switch (nt) {
    case 16: return launch_decode<16, kCG, OutT>(tw, tx0, tx1, p, workers, st);
    case 32: return launch_decode<32, kCG, OutT>(tw, tx0, tx1, p, workers, st);
    case 64: return launch_decode<64, kCG, OutT>(tw, tx0, tx1, p, workers, st);
    case 128: return launch_decode<128, kCG, OutT>(tw, tx0, tx1, p, workers, st);
    case 256: return launch_decode<256, kCG, OutT>(tw, tx0, tx1, p, workers, st);
    default: return SGLK_E_VARIANT;
  }
}

// Arguments were validated by gemm_entry (pointers, strides, alignment, dtype); M in [1, 256].
int decode_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M, int64_t N,
                 int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk, int64_t sb_sk,
                 int64_t sb_sn, int out_dtype, int cg, int num_sms, int flags, void* workspace, int64_t workspace_bytes,
                 cudaStream_t st) {
  if (num_sms <= 0) num_sms = device_sm_count();
  bool have_ws = workspace != nullptr && (reinterpret_cast<uintptr_t>(workspace) & 15) == 0;
  DecodePlan pl = make_decode_plan(M, N, K, cg, num_sms, have_ws, flags);
  if (pl.stream_k && workspace_bytes < kCounterAreaBytes + pl.partial_bytes) pl = make_decode_plan(M, N, K, cg, num_sms, false, flags);
  if (pl.stages < 2) return SGLK_E_SHAPE;

  DecodeParams p;
  memset(&p, 0, sizeof(p));
  p.sa = sa; p.sb = sb; p.C = Cout;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.sa_sm = sa_sm; p.sa_sk = sa_sk; p.sb_sk = sb_sk; p.sb_sn = sb_sn;
  p.num_tiles = pl.num_tiles;
  p.num_k_blocks = pl.num_kb;
  p.num_sb_cols = int32_t((N + 127) / 128);
  p.stream_k = pl.stream_k;
  p.sk_tile_base = pl.sk_tile_base;
  p.sk_workers = pl.sk_workers;
  p.num_stages = pl.stages;
  p.stage_bytes = pl.stage_bytes;
  p.scale_slot_bytes = pl.scale_slot;
  // activation rows per CTA: only rows that exist are loaded (TMA fills out-of-bounds box rows slowly), in whole
  // 8-row swizzle atoms; stale smem behind them only feeds accumulator columns that are never stored
  const int x_slot_rows = pl.nt / cg;
  for (int r = 0; r < 2; ++r) {
    int64_t rows = M - int64_t(r) * x_slot_rows;
    rows = rows < 0 ? 0 : (rows > x_slot_rows ? x_slot_rows : rows);
    p.x_rows[r] = (r < cg) ? uint32_t((rows + 7) / 8 * 8) : 0u;
  }
  p.stage_tx_bytes = uint32_t(cg) * kWTileBytes + (p.x_rows[0] + p.x_rows[1]) * BLOCK_K;
  p.ws_counters = pl.stream_k ? static_cast<int32_t*>(workspace) : nullptr;
  p.ws_partials = pl.stream_k ? reinterpret_cast<float*>(static_cast<char*>(workspace) + kCounterAreaBytes) : nullptr;

  CUtensorMap tw, tx0, tx1;
  if (!make_tmap_2d(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(N), uint64_t(K), uint64_t(ldb), BLOCK_M, BLOCK_K))
    return SGLK_E_DRIVER;
  if (!make_tmap_2d(&tx0, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, A, uint64_t(M), uint64_t(K), uint64_t(lda), p.x_rows[0], BLOCK_K))
    return SGLK_E_DRIVER;
  tx1 = tx0;
  if (p.x_rows[1] &&
      !make_tmap_2d(&tx1, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, A, uint64_t(M), uint64_t(K), uint64_t(lda), p.x_rows[1], BLOCK_K))
    return SGLK_E_DRIVER;
  if (cg == 2) {
    if (out_dtype == SGLK_BF16) return launch_decode_nt<2, __nv_bfloat16>(pl.nt, tw, tx0, tx1, p, pl.workers, st);
    return launch_decode_nt<2, __half>(pl.nt, tw, tx0, tx1, p, pl.workers, st);
  }
  if (out_dtype == SGLK_BF16) return launch_decode_nt<1, __nv_bfloat16>(pl.nt, tw, tx0, tx1, p, pl.workers, st);
  return launch_decode_nt<1, __half>(pl.nt, tw, tx0, tx1, p, pl.workers, st);
}

// ============================================================================================================
// Row-balanced swap-AB kernel (variant `swapab_rows`, M <= 128 tokens)
// ============================================================================================================
// The tile kernels above cut N into 128 / 256-row tiles and then have to balance a tile count that is not a multiple of
// the SM count (N=24576: 96 pair tiles on 74 pairs; N=2112: 17 tiles on 148 SMs) with a second, partly filled wave or
// with split-K partials and a fix-up.  A decode GEMM is a weight stream, and a stream balances by BYTES: here every CTA
// owns one contiguous range of R = ceil8(N / #SMs) weight rows (168 at N=24576, 16 at N=2112) for the WHOLE of K — no
// second wave, no partials, no workspace, no atomics.  A range is walked in passes of at most 256 rows; a pass is one or
// two MMA "sub-tiles" (128 rows + the rest), each with its own TMA box cut to the rows that exist (a box never reads
// more than 7 rows past the range; the MMA reads 128 smem rows regardless, the stale ones only feed TMEM lanes that are
// never stored) and its own accumulator columns in TMEM.  Rows of one range may lie in up to three 128-row weight-scale
// blocks, so sb is a per-LANE factor (lane = weight row), picked from four staged values.
// A ring stage holds G K blocks (4 for the small shapes: the issuer's and the promotion warps' per-hand-off latencies
// are paid once per 4 blocks) and is one accumulator hand-off.
struct RowsParams {
  const float* sa;
  const float* sb;
  void* C;
  int64_t M, N, K, ldc;
  int64_t sa_sm, sa_sk, sb_sk, sb_sn;
  int32_t num_k_blocks, num_sb_cols;
  int32_t rows_per_worker;   // R = passes * rows_per_pass
  int32_t passes;
  int32_t rows_per_pass;     // multiple of 8, <= 128 * S
  int32_t num_stages;
  // a ring stage holds G K blocks, operand-major: [sub-tile 0: G tiles][sub-tile 1: G tiles][activations: G tiles]; a tile
  // is box-rows x 128 bytes (box rows are multiples of 8, so every tile starts 1 KB aligned)
  uint32_t stage_bytes;
  uint32_t sub_tile_bytes[2], x_tile_bytes;
  uint32_t sub1_off, x_off;  // offsets of the sub-tile 1 / activation regions inside a stage
  uint32_t kb_tx_bytes;      // bytes TMA delivers per K block (boxes are always delivered whole, out-of-bounds zero-filled)
  int32_t kb3d;              // K % 128 == 0: the tensor maps are the [kblock][row][128 B] views, one TMA instruction per
                             // operand per stage; else 2-D maps (which zero-fill the K tail), one per operand per K block
  int32_t ksplit;            // 1, 2 or 4: CTAs of one CLUSTER share a row range and take one K range each; their partials
                             // are summed through distributed shared memory by the cluster's first CTA, in rank order
};

template <int NT, int S, int G>
struct RowsCfg {
  static constexpr int kEW = 8;                              // promotion / epilogue warps: two per TMEM lane quarter
  static constexpr int kThreads = 128 + kEW * 32;
  static constexpr int CW = NT / 2;                          // token columns per promotion warp
  static constexpr int IW = CW < 32 ? CW : 32;               // columns per tcgen05.ld item
  static constexpr int IPB = CW / IW;                        // items per (K block, sub-tile) per warp
  static constexpr int ITEMS = G * S * IPB;                  // items per hand-off
  static constexpr bool kOneBatch = ITEMS * IW <= 64;        // the whole hand-off fits the staging registers
  static constexpr int NBUF = kOneBatch ? ITEMS : 2;
  static constexpr int SLOT_COLS = G * S * NT;               // TMEM columns of one hand-off
  static constexpr int kSlots = (512 / SLOT_COLS) < 8 ? (512 / SLOT_COLS) : 8;
  static constexpr int TMEM_COLS = kSlots * SLOT_COLS;
  static constexpr int KB_SCALE_BYTES = (NT + 4) * 4;        // sa[0..NT) + four sb values per K block
  static constexpr int SCALE_SLOT = G * KB_SCALE_BYTES;      // per ring stage
  static_assert(NT % 16 == 0 && NT >= 16 && NT <= 128 && (S == 1 || S == 2), "shape");
  static_assert(kSlots >= 2 && (TMEM_COLS & (TMEM_COLS - 1)) == 0 && TMEM_COLS >= 32 && TMEM_COLS <= 512, "TMEM columns");
  static_assert(4 * 32 * 40 + kEW * 32 * 232 <= kThreads * 168, "register re-split");
};

constexpr int kRowsTailBytes = 16384;   // behind the ring: factor slots + barriers; also what the last sub-tile's MMA may over-read

template <int NT, int S, int G, typename OutT>
__global__ void __launch_bounds__(RowsCfg<NT, S, G>::kThreads, 1)
fp8_blockwise_gemm_rows_kernel(const __grid_constant__ CUtensorMap tmap_w0, const __grid_constant__ CUtensorMap tmap_w1,
                               const __grid_constant__ CUtensorMap tmap_x, const RowsParams p) {
  using C = RowsCfg<NT, S, G>;
  constexpr int CW = C::CW, kSlots = C::kSlots, kEW = C::kEW;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  if ((ptx::smem_u32(smem_raw) & 1023u) != 0) __trap();
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  const int nstages = p.num_stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t scale_base = smem_base + uint32_t(nstages) * stage_bytes;
  const uint32_t bar_base = scale_base + uint32_t(nstages) * C::SCALE_SLOT;          // 16-byte aligned
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kDecMaxStages + s); };
  auto sfull_bar = [&](int s) { return bar_base + 8u * (2 * kDecMaxStages + s); };
  auto sempty_bar = [&](int s) { return bar_base + 8u * (3 * kDecMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (4 * kDecMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (4 * kDecMaxStages + 8 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (4 * kDecMaxStages + 16);
  // (the host sized the ring so that nstages factor slots + the barriers fit the 16 KB tail: make_rows_plan)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_w0);
    if (S == 2) ptx::prefetch_tensormap(&tmap_w1);
    ptx::prefetch_tensormap(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < nstages; ++s) {
      ptx::mbar_init(full_bar(s), 1);                  // the producer's arrive.expect_tx
      ptx::mbar_init(empty_bar(s), 1);                 // one tcgen05.commit
      ptx::mbar_init(sfull_bar(s), 32);                // every lane of the scale loader
      ptx::mbar_init(sempty_bar(s), kEW);              // the promotion warps have read the slot
    }
    for (int s = 0; s < kSlots; ++s) {
      ptx::mbar_init(tfull_bar(s), 1);
      ptx::mbar_init(tempty_bar(s), kEW);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc<1>(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish<1>();
  }
  if (warp == 3) {
    // factor slots of tokens >= M are never filled: give them a defined value (their columns are never stored)
    for (uint32_t off = uint32_t(lane) * 4u; off < uint32_t(nstages) * C::SCALE_SLOT; off += 128u)
      *reinterpret_cast<float*>(smem_raw + (scale_base - smem_base) + off) = 0.0f;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_base));

  // this CTA's weight rows [row0, row_end), walked in `passes` passes of rows_per_pass rows, and its K blocks
  // [kb_begin, kb_begin + num_kb) (all of K unless the cluster splits it); every role runs the same loops
  const int ksplit = p.ksplit;
  const uint32_t kq = ksplit > 1 ? ptx::cluster_ctarank() : 0u;
  const int kb_begin = int(int64_t(p.num_k_blocks) * kq / ksplit);
  const int num_kb = int(int64_t(p.num_k_blocks) * (kq + 1) / ksplit) - kb_begin;
  const int num_groups = (num_kb + G - 1) / G;
  const int row0 = (int(blockIdx.x) / ksplit) * p.rows_per_worker;
  const int row_end = min(row0 + p.rows_per_worker, int(p.N));

  if (warp == 0) {
    // =========================== TMA producer ===========================
    ptx::reg_dealloc<40>();
    const uint64_t pol_w = ptx::policy_evict_first();      // weights pass through once
    const uint64_t pol_x = ptx::policy_evict_last();       // the activation K blocks are re-read by every CTA
    int stage = 0;
    uint32_t phase = 0;
    for (int pass = 0; pass < p.passes; ++pass) {
      const int base = row0 + pass * p.rows_per_pass;
      if (base >= row_end) break;
      for (int g = 0; g < num_groups; ++g) {
        const int nb = min(G, num_kb - g * G);
        ptx::mbar_wait(empty_bar(stage), phase ^ 1);
        if (ptx::elect_one()) {
          const uint32_t dst = smem_base + uint32_t(stage) * stage_bytes;
          if (p.kb3d) {
            // whole boxes of G K blocks (K blocks past this CTA's range — the next CTA's, or zero fill past K — are never multiplied)
            ptx::mbar_arrive_expect_tx(full_bar(stage), uint32_t(G) * p.kb_tx_bytes);
            ptx::tma_load_3d(&tmap_w0, full_bar(stage), dst, 0, base, kb_begin + g * G, pol_w);
            if constexpr (S == 2) ptx::tma_load_3d(&tmap_w1, full_bar(stage), dst + p.sub1_off, 0, base + BLOCK_M, kb_begin + g * G, pol_w);
            ptx::tma_load_3d(&tmap_x, full_bar(stage), dst + p.x_off, 0, 0, kb_begin + g * G, pol_x);
          } else {
            ptx::mbar_arrive_expect_tx(full_bar(stage), uint32_t(nb) * p.kb_tx_bytes);
            for (int j = 0; j < nb; ++j) {
              const int kb = kb_begin + g * G + j;
              ptx::tma_load_2d(&tmap_w0, full_bar(stage), dst + uint32_t(j) * p.sub_tile_bytes[0], kb * BLOCK_K, base, pol_w);
              if constexpr (S == 2)
                ptx::tma_load_2d(&tmap_w1, full_bar(stage), dst + p.sub1_off + uint32_t(j) * p.sub_tile_bytes[1], kb * BLOCK_K,
                                 base + BLOCK_M, pol_w);
              ptx::tma_load_2d(&tmap_x, full_bar(stage), dst + p.x_off + uint32_t(j) * p.x_tile_bytes, kb * BLOCK_K, 0, pol_x);
            }
          }
        }
        __syncwarp();
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
    }
    __syncwarp();
    if (ksplit > 1) ptx::cluster_sync();               // (1) the cluster's partials are in shared memory
  } else if (warp == 1 || warp == 2) {
    // =========================== MMA issuers ===========================
    // Two warps (the second one also owns the TMEM allocation) take alternate stages: at these widths the tensor pipe
    // needs 8-64 clk per MMA but one thread issues one only every ~35-40 clk (measured: ~175 clk per K block of four
    // MMAs, whatever the stage depth, load form or issue order), which bounded the small-N shapes.  MMAs of different
    // stages touch different smem and different TMEM columns, so the two issue streams need no ordering between them;
    // each thread's commits cover its own MMAs.
    ptx::reg_dealloc<40>();
    constexpr uint32_t idesc = make_idesc(BLOCK_M, NT);
    const uint32_t tmem_base_u = __shfl_sync(0xFFFFFFFFu, tmem_base, 0);
    const uint32_t desc_hi = uint32_t(make_smem_desc(0) >> 32);
    const uint32_t tile_lo[2] = {p.sub_tile_bytes[0] >> 4, p.sub_tile_bytes[1] >> 4};
    const uint32_t x_tile_lo = p.x_tile_bytes >> 4, sub1_lo = p.sub1_off >> 4, x_lo = p.x_off >> 4;
    int stage = 0, slot = 0, turn = warp - 1;          // this warp issues when turn == 0
    uint32_t phase = 0, slot_phase = 0;
    for (int pass = 0; pass < p.passes; ++pass) {
      if (row0 + pass * p.rows_per_pass >= row_end) break;
      for (int g = 0; g < num_groups; ++g) {
        if (turn == 0) {
          const int nb = min(G, num_kb - g * G);
          ptx::mbar_wait(full_bar(stage), phase);
          ptx::mbar_wait(tempty_bar(slot), slot_phase ^ 1);   // slot drained by every promotion warp?
          ptx::tc_fence_after();
          const uint32_t st_lo = ((smem_base + uint32_t(stage) * stage_bytes) & 0x3FFFFu) >> 4;
          const uint32_t d0 = tmem_base_u + uint32_t(slot * C::SLOT_COLS);
          if (ptx::elect_one()) {
            // The four K=32 MMAs of one (K block, sub-tile) accumulate into the same TMEM columns — a dependent chain;
            // the G x S chains of a stage are independent and are issued round-robin.
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
#pragma unroll
              for (int j = 0; j < G; ++j) {
                if (j < nb) {
                  const uint32_t b_lo = st_lo + x_lo + uint32_t(j) * x_tile_lo + uint32_t(k * (UMMA_K >> 4));
#pragma unroll
                  for (int s = 0; s < S; ++s) {
                    const uint32_t a_lo = st_lo + (s ? sub1_lo : 0u) + uint32_t(j) * tile_lo[s] + uint32_t(k * (UMMA_K >> 4));
                    dptx::mma_f8_lohi(d0 + uint32_t((j * S + s) * NT), a_lo, b_lo, desc_hi, idesc, k > 0 ? 1u : 0u);
                  }
                }
              }
            }
            ptx::tc_commit<1>(empty_bar(stage));     // smem stage reusable once these MMAs have read it
            ptx::tc_commit<1>(tfull_bar(slot));      // the hand-off's partials are complete
          }
          __syncwarp();
        }
        turn ^= 1;
        if (++stage == nstages) { stage = 0; phase ^= 1; }
        if (++slot == kSlots) { slot = 0; slot_phase ^= 1; }
      }
    }
    __syncwarp();
    if (ksplit > 1) ptx::cluster_sync();               // (1)
  } else if (warp == 3) {
    // =========================== scale loader ===========================
    // per K block: sa[token, kb] for every token and the sb[kb, .] of the (up to three) 128-row scale blocks this pass
    // touches -> the stage's factor slot, 4 bytes per cp.async (any strides), completion tracked by the slot's mbarrier
    ptx::reg_dealloc<40>();
    int stage = 0;
    uint32_t phase = 0;
    for (int pass = 0; pass < p.passes; ++pass) {
      const int base = row0 + pass * p.rows_per_pass;
      if (base >= row_end) break;
      const int sbc = min(base / 128 + min(lane, 3), p.num_sb_cols - 1);
      const float* sb_ptr = p.sb + int64_t(sbc) * p.sb_sn;
      for (int g = 0; g < num_groups; ++g) {
        const int nb = min(G, num_kb - g * G);
        ptx::mbar_wait(sempty_bar(stage), phase ^ 1);
        const uint32_t slot = scale_base + uint32_t(stage) * C::SCALE_SLOT;
        for (int j = 0; j < nb; ++j) {
          const int kb = kb_begin + g * G + j;
          const uint32_t kslot = slot + uint32_t(j) * C::KB_SCALE_BYTES;
#pragma unroll
          for (int q = 0; q < (NT + 31) / 32; ++q) {
            const int t = lane + 32 * q;
            if (t < NT && t < p.M) ptx::cp_async4(kslot + 4u * t, p.sa + int64_t(t) * p.sa_sm + int64_t(kb) * p.sa_sk);
          }
          if (lane < 4) ptx::cp_async4(kslot + 4u * NT + 4u * lane, sb_ptr + int64_t(kb) * p.sb_sk);
        }
        dptx::cp_async_mbar_arrive_noinc(sfull_bar(stage));
        if (++stage == nstages) { stage = 0; phase ^= 1; }
      }
    }
    ptx::cp_async_commit();
    ptx::cp_async_wait<0>();
    __syncwarp();
    if (ksplit > 1) ptx::cluster_sync();               // (1)
  } else {
    // =========================== promotion + epilogue warps ===========================
    ptx::reg_alloc<232>();
    bool synced = false;
    const int ew = warp - 4;
    const int quarter = warp & 3;                      // TMEM lanes this warp may touch: 32*quarter .. +31
    const int colgroup = ew >> 2;                      // which half of the token columns
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const int t0 = colgroup * CW;                      // first token column of this warp
    constexpr int IW = C::IW, IPB = C::IPB, ITEMS = C::ITEMS, NBUF = C::NBUF;
    int stage = 0, slot = 0;
    uint32_t phase = 0, slot_phase = 0;
    OutT* Cp = reinterpret_cast<OutT*>(p.C);
    const bool pair_ok = (p.ldc % 2 == 0) && (p.N % 2 == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 3) == 0);
    for (int pass = 0; pass < p.passes; ++pass) {
      const int base = row0 + pass * p.rows_per_pass;
      if (base >= row_end) break;
      const int limit = min(base + p.rows_per_pass, row_end);       // rows [base, limit) are this pass's output columns
      // byte offset (inside a K block's factor slot) of this lane's weight-scale value, per sub-tile
      uint32_t sb_off[S];
#pragma unroll
      for (int s = 0; s < S; ++s) {
        const int n = base + s * BLOCK_M + quarter * 32 + lane;
        sb_off[s] = 4u * uint32_t(NT) + 4u * uint32_t(min(n / 128 - base / 128, 3));
      }
      float accum[S][CW];
#pragma unroll
      for (int s = 0; s < S; ++s)
#pragma unroll
        for (int i = 0; i < CW; ++i) accum[s][i] = 0.0f;
      for (int g = 0; g < num_groups; ++g) {
        const int nb = min(G, num_kb - g * G);
        ptx::mbar_wait(tfull_bar(slot), slot_phase);   // the hand-off's MMAs complete
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + lane_base + uint32_t(slot * C::SLOT_COLS + t0);
        const uint32_t fslot = scale_base + uint32_t(stage) * C::SCALE_SLOT;
        uint32_t v[NBUF][IW];
        // item i = (K block j, sub-tile s, column chunk c) with i = (j * S + s) * IPB + c
        auto item_ld = [&](int i, uint32_t (&buf)[IW]) {
          dptx::tmem_ld<IW>(taddr + uint32_t((i / IPB) * NT + (i % IPB) * IW), buf);
        };
        auto item_fma = [&](int i, const uint32_t (&buf)[IW]) {
          const int j = i / (S * IPB), s = (i / IPB) % S, c = i % IPB;
          const uint32_t kslot = fslot + uint32_t(j) * C::KB_SCALE_BYTES;
          const float sbv = ptx::lds_f32(kslot + sb_off[s]);
          const uint32_t fac = kslot + 4u * uint32_t(t0 + c * IW);
#pragma unroll
          for (int q = 0; q < IW / 4; ++q) {
            const float4 f = dptx::lds_f32x4(fac + 16u * q);
            float* a = &accum[s][c * IW + 4 * q];
            dptx::fma2_scaled(a[0], a[1], buf[4 * q], buf[4 * q + 1], f.x, f.y, sbv);
            dptx::fma2_scaled(a[2], a[3], buf[4 * q + 2], buf[4 * q + 3], f.z, f.w, sbv);
          }
        };
        auto release_slot = [&]() {                    // every tcgen05.ld of this warp on the slot has completed
          ptx::tc_fence_before();
          if (lane == 0) ptx::mbar_arrive(tempty_bar(slot));
        };
        if constexpr (C::kOneBatch) {
          // the whole hand-off in one go: all tcgen05.ld issued, one wait, slot handed back, then the FMAs
#pragma unroll
          for (int i = 0; i < ITEMS; ++i)
            if (i / (S * IPB) < nb) item_ld(i, v[i]);
          ptx::mbar_wait(sfull_bar(stage), phase);     // the factors of these K blocks are in smem (long since, normally)
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          release_slot();
#pragma unroll
          for (int i = 0; i < ITEMS; ++i)
#pragma unroll
            for (int r = 0; r < IW; ++r) asm volatile("" : "+r"(v[i][r]));     // consumers stay below the wait
          // per (K block, column chunk): the activation factors are read from smem ONCE and applied to both sub-tiles
#pragma unroll
          for (int j = 0; j < G; ++j) {
            if (j < nb) {
              const uint32_t kslot = fslot + uint32_t(j) * C::KB_SCALE_BYTES;
              float sbv[S];
#pragma unroll
              for (int s = 0; s < S; ++s) sbv[s] = ptx::lds_f32(kslot + sb_off[s]);
#pragma unroll
              for (int c = 0; c < IPB; ++c) {
                const uint32_t fac = kslot + 4u * uint32_t(t0 + c * IW);
#pragma unroll
                for (int q = 0; q < IW / 4; ++q) {
                  const float4 f = dptx::lds_f32x4(fac + 16u * q);
#pragma unroll
                  for (int s = 0; s < S; ++s) {
                    const uint32_t* buf = v[(j * S + s) * IPB + c];
                    float* a = &accum[s][c * IW + 4 * q];
                    dptx::fma2_scaled(a[0], a[1], buf[4 * q], buf[4 * q + 1], f.x, f.y, sbv[s]);
                    dptx::fma2_scaled(a[2], a[3], buf[4 * q + 2], buf[4 * q + 3], f.z, f.w, sbv[s]);
                  }
                }
              }
            }
          }
        } else {
          // item i+1 is in flight while item i is folded in; the slot goes back as soon as its last item is in registers
          ptx::mbar_wait(sfull_bar(stage), phase);
          const int last = nb * S * IPB - 1;
          item_ld(0, v[0]);
#pragma unroll
          for (int i = 0; i < ITEMS; ++i) {
            if (i <= last) {
              dptx::tmem_ld_wait<IW>(v[i & 1]);
              if (i < last) item_ld(i + 1, v[(i + 1) & 1]);
              else release_slot();
              item_fma(i, v[i & 1]);
            }
          }
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(sempty_bar(stage));     // every factor of this stage has been read
        if (++stage == nstages) { stage = 0; phase ^= 1; }
        if (++slot == kSlots) { slot = 0; slot_phase ^= 1; }
      }
      // ---------------- split K: the cluster's partials meet in the first CTA ----------------
      if (ksplit > 1) {
        // (one pass per CTA in this mode.)  Every MMA of this CTA has completed — this warp has seen the commit of
        // every hand-off — so the operand ring is free and holds the partial: [warp][value][lane] as float4.
        constexpr int V4 = S * CW / 4;
        if (kq != 0) {
          float4* red = reinterpret_cast<float4*>(smem_raw) + ew * V4 * 32 + lane;
#pragma unroll
          for (int s = 0; s < S; ++s)
#pragma unroll
            for (int j = 0; j < CW / 4; ++j)
              red[(s * (CW / 4) + j) * 32] = make_float4(accum[s][4 * j], accum[s][4 * j + 1], accum[s][4 * j + 2], accum[s][4 * j + 3]);
        }
        ptx::cluster_sync();                           // (1): release / acquire at cluster scope
        synced = true;
        if (kq != 0) break;
        // fixed order (bit-reproducible): own + rank 1 + rank 2 + ...
        for (uint32_t q = 1; q < uint32_t(ksplit); ++q) {
          const uint32_t peer = ptx::mapa(smem_base, q) + uint32_t(ew * V4 * 32 + lane) * 16u;
#pragma unroll
          for (int s = 0; s < S; ++s)
#pragma unroll
            for (int j = 0; j < CW / 4; ++j) {
              float4 t;
              asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
                           : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                           : "r"(peer + uint32_t((s * (CW / 4) + j) * 32) * 16u)
                           : "memory");
              accum[s][4 * j] += t.x; accum[s][4 * j + 1] += t.y; accum[s][4 * j + 2] += t.z; accum[s][4 * j + 3] += t.w;
            }
        }
      }
      // ---------------- C^T out of the registers: lane = output column n, accum[s][j] = token t0 + j ----------------
#pragma unroll
      for (int s = 0; s < S; ++s) {
        const int n = base + s * BLOCK_M + quarter * 32 + lane;
        if (pair_ok) {
          // neighbouring lanes swap one value each: the even lane ends up with (n, n+1) of token j, the odd lane with
          // (n-1, n) of token j+1 — one 4-byte store per two values (base and limit are even: a pair is never split)
          const int odd = lane & 1;
          const int col = n - odd;
#pragma unroll
          for (int j = 0; j < CW; j += 2) {
            const float send = odd ? accum[s][j] : accum[s][j + 1];
            const float recv = __shfl_xor_sync(0xFFFFFFFFu, send, 1);
            const uint32_t w = odd ? pack2<OutT>(recv, accum[s][j + 1]) : pack2<OutT>(accum[s][j], recv);
            const int tok = t0 + j + odd;
            if (tok < p.M && col < limit) *reinterpret_cast<uint32_t*>(Cp + int64_t(tok) * p.ldc + col) = w;
          }
        } else if (n < limit) {
#pragma unroll
          for (int j = 0; j < CW; ++j) {
            if (t0 + j < p.M) {
              const uint32_t w = pack2<OutT>(accum[s][j], 0.0f);
              Cp[int64_t(t0 + j) * p.ldc + n] = *reinterpret_cast<const OutT*>(&w);
            }
          }
        }
      }
    }
    __syncwarp();
    if (ksplit > 1 && !synced) ptx::cluster_sync();    // (1) — a CTA without rows (the host never launches one)
  }

  // =========================== teardown ===========================
  ptx::tc_fence_before();
  if (ksplit > 1) ptx::cluster_sync();                 // (2) nobody leaves while the first CTA may still read its shared memory
  else __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<1>(tmem_base, C::TMEM_COLS);
}

struct RowsPlan {
  int nt, S, G, ksplit, workers, rows_per_worker, passes, rows_per_pass, stages;
  uint32_t sub_rows[2], x_rows, stage_bytes, sub1_off, x_off, kb_tx_bytes;
};

bool rows_supported(int64_t M, int64_t N, int64_t K) { return M >= 1 && M <= 128 && N >= 1 && K >= 1; }

static int rows_max_clusters(int cs, int num_sms);

static RowsPlan make_rows_plan(int64_t M, int64_t N, int64_t K, int num_sms, int flags) {
  RowsPlan pl;
  pl.nt = decode_nt(M);
  const int64_t W = num_sms < 1 ? 1 : num_sms;
  const int64_t num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  // K split over the CTAs of a cluster.  An MMA reads its whole 128-row A tile from shared memory whatever part of it
  // holds rows of the range (measured: ~70 clk per K=32 MMA at these widths, i.e. ~0.14 us per K block per CTA), so a
  // narrow weight spread over all SMs as 16-row ranges is bound by 56 K blocks of mostly empty MMAs (N=2112: 11.4 us).
  // Splitting K four ways gives every CTA a quarter of the K blocks on four times the rows, and a quarter of the
  // activation re-reads.  Used while there are fewer 128-row tiles than clusters; flags bits 16..23 override (1, 2, 4).
  int ks = (flags >> 16) & 0xFF;
  if (ks == 0) {
    const int64_t tiles128 = (N + 127) / 128;
    ks = (num_kb >= 8 && tiles128 * 4 <= W) ? 4 : ((num_kb >= 8 && tiles128 * 2 <= W) ? 2 : 1);
  }
  if (ks != 1 && ks != 2 && ks != 4) ks = 1;
  if (ks > num_kb) ks = 1;
  int64_t groups = W;
  if (ks > 1) {
    groups = rows_max_clusters(ks, int(W));
    if (groups < 1) { ks = 1; groups = W; }
  }
  int64_t R = ((N + groups - 1) / groups + 7) / 8 * 8;        // rows per CTA (cluster), whole 8-row swizzle atoms
  if (ks > 1 && R > 256) {                                    // split K runs one pass per CTA
    ks = 1; groups = W;
    R = ((N + groups - 1) / groups + 7) / 8 * 8;
  }
  pl.ksplit = ks;
  pl.passes = int((R + 255) / 256);
  pl.rows_per_pass = int(((R + pl.passes - 1) / pl.passes + 7) / 8 * 8);
  R = int64_t(pl.rows_per_pass) * pl.passes;
  pl.rows_per_worker = int(R);
  pl.workers = int((N + R - 1) / R);
  pl.S = pl.rows_per_pass > BLOCK_M ? 2 : 1;
  pl.sub_rows[0] = uint32_t(pl.rows_per_pass > BLOCK_M ? BLOCK_M : pl.rows_per_pass);
  pl.sub_rows[1] = pl.S == 2 ? uint32_t(pl.rows_per_pass - BLOCK_M) : 0u;
  pl.x_rows = uint32_t((M + 7) / 8 * 8);
  // K blocks per ring stage / accumulator hand-off: more of them = more independent MMA chains in flight and fewer
  // hand-offs; bounded by TMEM (a hand-off is G x S x NT columns and two must fit) and by the ring (three stages at
  // least, or the loads cannot run ahead of the MMAs).  Candidates are the instantiations launch_rows_cfg knows.
  pl.kb_tx_bytes = (pl.sub_rows[0] + pl.sub_rows[1] + pl.x_rows) * BLOCK_K;
  {
    const int first = pl.S == 2 ? (pl.nt <= 32 ? 2 : 1) : (pl.nt == 16 ? 8 : (pl.nt <= 64 ? 4 : 2));
    const int second = pl.S == 2 ? first : (pl.nt == 128 ? 1 : 2);
    pl.G = (kMaxDynSmem - kRowsTailBytes) / int(uint32_t(first) * pl.kb_tx_bytes) >= 3 ? first : second;
  }
  // tiles are box-rows x 128 B (multiples of 1 KB); the MMA reads 128 weight rows / NT activation rows from a tile's base
  // whatever the box held: what lies behind (the next tile, the next stage, the 16 KB tail) is addressable smem
  pl.sub1_off = uint32_t(pl.G) * pl.sub_rows[0] * BLOCK_K;
  pl.x_off = pl.sub1_off + uint32_t(pl.G) * pl.sub_rows[1] * BLOCK_K;
  pl.stage_bytes = uint32_t(pl.G) * pl.kb_tx_bytes;
  int n = (kMaxDynSmem - kRowsTailBytes) / int(pl.stage_bytes);
  if (n > kDecMaxStages) n = kDecMaxStages;
  const int scale_slot = pl.G * (pl.nt + 4) * 4, bar_bytes = 8 * (4 * kDecMaxStages + 16) + 16;
  while (n > 2 && n * scale_slot + bar_bytes > kRowsTailBytes) --n;
  pl.stages = n;
  return pl;
}

template <int NT, int S, int G, typename OutT>
static int launch_rows(const CUtensorMap& tw0, const CUtensorMap& tw1, const CUtensorMap& tx, const RowsParams& p,
                       int workers, cudaStream_t st) {
  auto kern = fp8_blockwise_gemm_rows_kernel<NT, S, G, OutT>;
  static std::atomic<uint64_t> configured{0};
  if (int e = ensure_max_dyn_smem(kern, configured, kMaxDynSmem)) return e;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(unsigned(workers * p.ksplit));
  cfg.blockDim = dim3(RowsCfg<NT, S, G>::kThreads);
  cfg.dynamicSmemBytes = kMaxDynSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = unsigned(p.ksplit);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return int(cudaLaunchKernelEx(&cfg, kern, tw0, tw1, tx, p));
}

// How many clusters of `cs` CTAs of this kernel (one CTA per SM: 227 KB of shared memory) the current device runs at
// once — clusters are placed inside a GPC, so it is less than #SMs / cs.  Every instantiation has the same footprint;
// one is asked, once per (device, cluster size).
static int rows_max_clusters(int cs, int num_sms) {
  static std::atomic<int> cache[64][5];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || cs < 1 || cs > 4) return num_sms / (cs < 1 ? 1 : cs);
  int n = cache[dev][cs].load(std::memory_order_acquire);
  if (n == 0) {
    auto kern = fp8_blockwise_gemm_rows_kernel<16, 1, 8, __nv_bfloat16>;
    n = -1;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem) == cudaSuccess) {
      cudaLaunchConfig_t cfg;
      memset(&cfg, 0, sizeof(cfg));
      cfg.gridDim = dim3(unsigned(cs * 64));
      cfg.blockDim = dim3(RowsCfg<16, 1, 8>::kThreads);
      cfg.dynamicSmemBytes = kMaxDynSmem;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = unsigned(cs);
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int got = 0;
      if (cudaOccupancyMaxActiveClusters(&got, kern, &cfg) == cudaSuccess && got > 0) n = got;
    }
    (void)cudaGetLastError();
    cache[dev][cs].store(n, std::memory_order_release);
  }
  if (n < 0) return 0;
  const int cap = num_sms / cs;                       // a caller that restricts num_sms restricts the clusters too
  return n < cap ? n : cap;
}

template <typename OutT>
static int launch_rows_cfg(const RowsPlan& pl, const CUtensorMap& tw0, const CUtensorMap& tw1, const CUtensorMap& tx,
                           const RowsParams& p, cudaStream_t st) {
  const int key = pl.nt * 100 + pl.S * 10 + pl.G;
  switch (key) {
    case 1618: return launch_rows<16, 1, 8, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 1612: return launch_rows<16, 1, 2, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 1622: return launch_rows<16, 2, 2, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 3214: return launch_rows<32, 1, 4, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 3212: return launch_rows<32, 1, 2, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 3222: return launch_rows<32, 2, 2, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 6414: return launch_rows<64, 1, 4, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 6412: return launch_rows<64, 1, 2, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 6421: return launch_rows<64, 2, 1, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 12812: return launch_rows<128, 1, 2, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 12811: return launch_rows<128, 1, 1, OutT>(tw0, tw1, tx, p, pl.workers, st);
    case 12821: return launch_rows<128, 2, 1, OutT>(tw0, tw1, tx, p, pl.workers, st);
    default: return SGLK_E_VARIANT;
  }
}

// Arguments were validated by gemm_entry (pointers, strides, alignment, dtype); M in [1, 128].
int rows_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M, int64_t N, int64_t K,
               int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk, int64_t sb_sk, int64_t sb_sn,
               int out_dtype, int num_sms, int flags, cudaStream_t st) {
  if (num_sms <= 0) num_sms = device_sm_count();
  const RowsPlan pl = make_rows_plan(M, N, K, num_sms, flags);
  if (pl.stages < 2) return SGLK_E_SHAPE;
  // the split-K partial of a CTA (S * NT/2 KB) lives in the operand ring
  if (pl.ksplit > 1 && int64_t(pl.S) * (pl.nt / 2) * 1024 > int64_t(pl.stages) * pl.stage_bytes) return SGLK_E_SHAPE;
  RowsParams p;
  memset(&p, 0, sizeof(p));
  p.sa = sa; p.sb = sb; p.C = Cout;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.sa_sm = sa_sm; p.sa_sk = sa_sk; p.sb_sk = sb_sk; p.sb_sn = sb_sn;
  p.num_k_blocks = int32_t((K + BLOCK_K - 1) / BLOCK_K);
  p.num_sb_cols = int32_t((N + 127) / 128);
  p.rows_per_worker = pl.rows_per_worker;
  p.passes = pl.passes;
  p.rows_per_pass = pl.rows_per_pass;
  p.num_stages = pl.stages;
  p.stage_bytes = pl.stage_bytes;
  p.sub_tile_bytes[0] = pl.sub_rows[0] * BLOCK_K;
  p.sub_tile_bytes[1] = pl.sub_rows[1] * BLOCK_K;
  p.x_tile_bytes = pl.x_rows * BLOCK_K;
  p.ksplit = pl.ksplit;
  p.kb3d = (K % BLOCK_K == 0 && ldb % 16 == 0 && lda % 16 == 0 && !(flags & SGLK_GEMM_DECODE_2D_LOADS)) ? 1 : 0;
  p.sub1_off = pl.sub1_off;
  p.x_off = pl.x_off;
  p.kb_tx_bytes = pl.kb_tx_bytes;
  CUtensorMap tw0, tw1, tx;
  if (p.kb3d) {
    // (an encode failure here is not an error: the 2-D maps below serve every shape)
    bool ok = make_tmap_kblocks(&tw0, B, uint64_t(N), uint64_t(K), uint64_t(ldb), pl.sub_rows[0], uint32_t(pl.G));
    tw1 = tw0;
    if (ok && pl.S == 2) ok = make_tmap_kblocks(&tw1, B, uint64_t(N), uint64_t(K), uint64_t(ldb), pl.sub_rows[1], uint32_t(pl.G));
    if (ok) ok = make_tmap_kblocks(&tx, A, uint64_t(M), uint64_t(K), uint64_t(lda), pl.x_rows, uint32_t(pl.G));
    if (!ok) p.kb3d = 0;
  }
  if (!p.kb3d) {
    if (!make_tmap_2d(&tw0, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(N), uint64_t(K), uint64_t(ldb), pl.sub_rows[0], BLOCK_K))
      return SGLK_E_DRIVER;
    tw1 = tw0;
    if (pl.S == 2 &&
        !make_tmap_2d(&tw1, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(N), uint64_t(K), uint64_t(ldb), pl.sub_rows[1], BLOCK_K))
      return SGLK_E_DRIVER;
    if (!make_tmap_2d(&tx, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, A, uint64_t(M), uint64_t(K), uint64_t(lda), pl.x_rows, BLOCK_K))
      return SGLK_E_DRIVER;
  }
  if (out_dtype == SGLK_BF16) return launch_rows_cfg<__nv_bfloat16>(pl, tw0, tw1, tx, p, st);
  return launch_rows_cfg<__half>(pl, tw0, tw1, tx, p, st);
}

}  // namespace sglk
