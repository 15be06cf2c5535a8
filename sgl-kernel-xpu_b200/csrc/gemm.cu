// fp8 (e4m3) blockwise-scaled GEMM for sm_100a:  C = sum_kb (sa[:,kb] x sb[kb,:]) .* (A_kb @ B_kb)
//
// Implements the op the reference only declares (include/sgl_kernel_ops.h:581-586,
// python/sgl_kernel/gemm.py:24-31); semantics from tests/test_fp8_blockwise_gemm.py:23-63.
//
// Design (B200-first; nothing here derives from the reference, which ships no kernel for this op):
//   * persistent, warp-specialised CTA: warp 0 = TMA producer, warp 1 = tcgen05.mma issuer,
//     warp 2 = TMEM allocator, warps 4..11 = promotion/epilogue warps.
//   * operands are both K-major e4m3, staged by TMA (128-byte swizzle) into a kStages-deep
//     shared-memory ring; one stage = one 128-deep K block = one scale block.
//   * fp32 (not UE8M0) scales rule out the hardware block-scaled MMA kinds, so every K block is
//     accumulated on its own into a TMEM accumulator stage (tcgen05.mma.kind::f8f6f4, 4 x K=32,
//     first one overwrites) and *promoted* in registers: the epilogue warps tcgen05.ld the
//     128 x BLOCK_N partial, and do acc += partial * (sa[m,kb]*sb[kb,nb]) with packed FFMA2.
//     kAccStages TMEM stages let the MMA of block kb+1.. run under the promotion of block kb.
//   * kCG == 2 pairs two CTAs (cta_group::2): one 256 x BLOCK_N tile per pair, each CTA loads its
//     own 128 rows of A and half of the B tile; the leader CTA issues the MMAs for both.
//   * C leaves through per-warp swizzled smem boxes and TMA stores (coalesced, clipped at M/N
//     edges by the tensor map); a plain predicated store path covers unaligned ldc.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "gemm_common.cuh"
#include "quant_kernel.cuh"   // div_pair / silu_mul: the fused output quantisation shares the quant op's arithmetic

namespace sglk {

constexpr int kNumEpiWarps = 8;
constexpr int kNumThreads = 128 + kNumEpiWarps * 32;   // 4 control warps + 8 epilogue warps
constexpr int kStoreBoxCols = 64;                      // bf16 columns per TMA store box (128 B)
constexpr int kMaxStages = 16;                         // barrier slots reserved for the smem ring
constexpr int kSmemOverhead = 1024 + 2048;             // barriers + sb slots (1 KB), per-thread sa slots (2 KB); base is 1024-aligned

struct GemmParams {
  const float* sa;
  const float* sb;
  void* C;
  int64_t M, N, K;
  int64_t ldc;
  int64_t sa_sm, sa_sk, sb_sk, sb_sn;
  int32_t num_m_blocks, num_n_blocks, num_k_blocks, num_sb_cols;
  int32_t group_m;   // tile rasterisation: tiles of `group_m` consecutive m-blocks share B panels in L2
  // stream-K (decode shapes / ragged waves): the (tile, K-block) units are cut into one contiguous
  // range per worker; tiles shared by several workers are summed through `ws_partials` by whichever
  // worker arrives last (fixed summation order -> deterministic)
  int32_t stream_k;
  int32_t scale_vec;         // scales contiguous along K blocks and 16-byte aligned: vector-load path of the promotion warps
  int32_t num_dests;         // TMA-store destinations (1, or the group size for the fused all-gather)
  // fused output quantisation (kEpi != 0): C is not written; e4m3 codes leave through the TMA-store map, one fp32
  // scale per (row, 128-column group) goes to q_scales[row * qs_sm + group * qs_sg]
  float* q_scales;
  int64_t qs_sm, qs_sg;
  float q_eps;
  long long* dbg;            // diagnostics (kDiag == 4): per-K-block clock64 stamps of CTA 0, else null
  void* mc_ptr;              // NVLS multicast address of this rank's panel (fused all-gather, multicast form) or null
  uint32_t stage_tx_bytes;   // bytes TMA delivers per stage to the (leader's) full barrier
  uint32_t stage_bytes;      // smem bytes of one ring stage = a_slot_bytes + B tile bytes (multiple of 1024)
  uint32_t a_slot_bytes;     // smem reserved for the A tile of a stage (rows that exist, rounded to 1 KB)
  int32_t num_stages;
  int32_t* ws_counters;   // [num_tiles][kCG*8], zero on entry, reset by the reducer
  float* ws_partials;     // [2*workers][kCG*8][NC/4][32] float4
  // grouped (MoE) mode, kGrouped instantiation only: A / C / sa are the experts' row ranges stacked, B is [E*N, K]
  const int4* grp_tiles;          // one entry per 128-row m-tile: {first row in the stack, valid rows, expert, 0}
  const int32_t* grp_num_m_tiles; // how many entries (written on the device by build_group_tiles_kernel)
  int64_t grp_sb_expert_stride;   // floats between the experts' weight-scale matrices
  // gated mode (kWaitA instantiation, sglk_fp8_blockwise_scaled_mm_gated): the rows of A (and of sa) arrive while the
  // kernel runs — written by the row-sharded quantisation of every rank (quant_bcast.cu) — one arrival counter per
  // 256-row m-block; a tile is started once int32(counter - a_ready_target) >= 0.  The m-blocks are walked own panel
  // first, then one "slice" (the j-th m-block of every other rank's panel, nearest rank first) after the other: the
  // order in which the panels, all sent at once, fill up.
  const uint32_t* a_ready;
  uint32_t a_ready_target;
  uint32_t a_wait_polls;          // poll rounds before a waiting CTA gives up with __trap(); 0 = wait for ever
  int32_t panel_own;              // m-blocks per rank panel
  int32_t panel_rank, panel_world;
  int32_t panel_group;            // remote m-blocks per rasterisation group (a whole number of slices)
  int32_t panel_whole;            // 1: the other panels arrive one after the other (nearest rank first) instead of slice by slice
};

constexpr int kMaxDests = 8;
struct CMaps { CUtensorMap m[kMaxDests]; };   // output tensor maps: [0] = local C, [1..] = peers (fused all-gather)

// One worker's walk over its share of the problem: whole tiles round-robin, or a contiguous range of
// (tile, K-block) units.  Producer, MMA issuer and epilogue warps all run the same iterator.
struct WorkIter {
  int64_t cur, end;
  int step, num_kb;
  bool stream_k;
  __device__ WorkIter(const GemmParams& p, int worker, int num_workers, int num_tiles) {
    num_kb = p.num_k_blocks;
    stream_k = p.stream_k != 0;
    if (stream_k) {
      const int64_t total = int64_t(num_tiles) * num_kb;
      cur = total * worker / num_workers;
      end = total * (worker + 1) / num_workers;
      step = 0;
    } else {
      cur = worker; end = num_tiles; step = num_workers;
    }
  }
  // next segment: tile index and K-block range [kb0, kb1); `first` = first segment of this worker
  __device__ bool next(int& tile, int& kb0, int& kb1) {
    if (cur >= end) return false;
    if (stream_k) {
      tile = int(cur / num_kb);
      kb0 = int(cur - int64_t(tile) * num_kb);
      const int64_t left = end - cur;
      kb1 = left < int64_t(num_kb - kb0) ? kb0 + int(left) : num_kb;
      cur += kb1 - kb0;
    } else {
      tile = int(cur); kb0 = 0; kb1 = num_kb;
      cur += step;
    }
    return true;
  }
};

template <int BLOCK_N, int kCG>
struct Cfg {
  static constexpr int LOAD_N = BLOCK_N / kCG;                 // B rows this CTA loads per stage
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K;            // 16 KB
  static constexpr int B_BYTES = LOAD_N * BLOCK_K;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int NC = BLOCK_N / 2;                       // columns per epilogue warp
  static constexpr int kAccStages = (512 / BLOCK_N) < 4 ? (512 / BLOCK_N) : 4;
  static constexpr int TMEM_COLS = kAccStages * BLOCK_N;       // 256 or 512: power of two
  static constexpr int STORE_BYTES = kNumEpiWarps * 32 * kStoreBoxCols * 2;   // 32 KB
  static constexpr int BAR_BYTES = 1024;
  static constexpr int smem_bytes(int stages, bool tma_store) {
    return 1024 + stages * STAGE_BYTES + (tma_store ? STORE_BYTES : 0) + BAR_BYTES;
  }
  static_assert(BLOCK_N % 64 == 0 && BLOCK_N <= 256, "BLOCK_N");
  static constexpr bool kCanTmaStore = (NC % kStoreBoxCols) == 0;
  static_assert((TMEM_COLS & (TMEM_COLS - 1)) == 0 && TMEM_COLS >= 32 && TMEM_COLS <= 512, "TMEM columns");
};

// tile index -> (m_block, n_block): groups of `group_m` m-blocks are walked n-major so that the
// CTAs resident at one time share a few B panels and `group_m` A panels in L2.
__device__ __forceinline__ void tile_coords(int tile, const GemmParams& p, int& mb, int& nb) {
  const int per_group = p.group_m * p.num_n_blocks;
  const int g = tile / per_group;
  const int first_m = g * p.group_m;
  const int rows = min(p.group_m, p.num_m_blocks - first_m);
  const int in_g = tile - g * per_group;
  mb = first_m + in_g % rows;
  nb = in_g / rows;
}

// gated mode: tile index -> (physical m-block, n-block) in arrival order (see GemmParams::a_ready)
__host__ __device__ __forceinline__ void tile_coords_panels(int tile, const GemmParams& p, int& mb, int& nb) {
  // straight-line on purpose (selects, no branches): with branches here ptxas kept the producer's loop state out of the
  // uniform registers
  const int own = p.panel_own, nn = p.num_n_blocks;
  const bool is_own = tile < own * nn;
  const int t = is_own ? tile : tile - own * nn;                  // own panel: groups of group_m m-blocks, n-major;
  const int gsz = is_own ? p.group_m : p.panel_group;             // the other panels: groups of panel_group
  const int span = is_own ? own : p.num_m_blocks - own;
  const int per_group = gsz * nn;
  const int g = t / per_group, first = g * gsz;
  const int rows = gsz < span - first ? gsz : span - first, in_g = t - g * per_group;
  nb = in_g / rows;
  const int r = first + (in_g - nb * rows);                       // position inside the own / the remote part of the walk
  // remote part: whole panels one after the other (i = peer index, j = block in panel), or slice by slice
  const int d = p.panel_whole ? own : (p.panel_world > 2 ? p.panel_world - 1 : 1);
  const int q = r / d, m = r - q * d;
  const int i = p.panel_whole ? q : m, j = p.panel_whole ? m : q;
  int panel = p.panel_rank + 1 + i;
  panel -= panel >= p.panel_world ? p.panel_world : 0;
  mb = is_own ? p.panel_rank * own + r : panel * own + j;
}
// kDiag (diagnostics only, results are NOT the op's): 1 = skip the promotion arithmetic, 2 = accumulate all of K in
// TMEM like an unscaled GEMM (one accumulator hand-off per tile) — both measure ceilings of the TMA/MMA pipeline.
// kSplitN: each K block is issued as two N/2-wide MMA halves with their own full/empty barriers, so a half can be
// promoted (and handed back) while the other half and the next K block are still in the tensor pipe.
// kEpi: 0 = C in OutT; 1 = fused per-token-group quantisation of the OutT-rounded result (e4m3 codes + one fp32 scale per
// 1x128 group, bit-identical to GEMM -> sgl_per_token_group_quant_8bit); 2 = the same after silu(gate)*up, the tile's
// first 128 columns being a gate block and its last 128 the matching up block (weights interleaved by 128-row blocks).
template <int BLOCK_N, int kStages, int kCG, typename OutT, bool kTmaStore, int kDiag = 0, bool kSplitN = false,
          bool kGrouped = false, int kEpi = 0, bool kMcB = false, int kPipe = 0, bool kWaitA = false>
__global__ void __launch_bounds__(kNumThreads, 1)
fp8_blockwise_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                          const __grid_constant__ CMaps tmap_c, const GemmParams p) {
  using C = Cfg<BLOCK_N, kCG>;
  // kMcB: a 2-CTA cluster of INDEPENDENT cta_group::1 tiles (same n-block, adjacent m-blocks).  Each CTA loads its own
  // A rows and HALF of the shared B tile, multicast into both CTAs' smem — the L2 traffic of the paired 256x256 tile,
  // but every accumulator hand-off (tcgen05.commit -> promotion warps -> issuer) stays inside one CTA.
  constexpr int kCl = kMcB ? 2 : kCG;                  // CTAs per cluster = per 256-row (kCl == 2) tile slot
  static_assert(!kMcB || (kCG == 1 && !kSplitN && !kGrouped && kDiag == 0), "multicast-B mode: cta_group::1 MMAs");
  static_assert(!kWaitA || (kCG == 2 && !kGrouped && kDiag == 0 && kEpi == 0 && !kMcB), "gated mode: the paired prefill tiles");
  constexpr int NC = C::NC;
  constexpr int kAccStages = C::kAccStages;
  constexpr int kHalves = kSplitN ? 2 : 1;             // accumulator hand-off granularity: whole stage or column half
  constexpr int kAccBars = kAccStages * kHalves;
  static_assert(!kSplitN || (BLOCK_N == 256 && (kDiag == 0 || kDiag == 4)), "split-N is for the 256-wide product tiles");
  static_assert(!kTmaStore || C::kCanTmaStore, "TMA-store epilogue needs >= 64 columns per warp");
  static_assert(!kGrouped || (kDiag == 0 && ((kCG == 1 && !kTmaStore && !kSplitN) || (kCG == 2 && kTmaStore && kSplitN))),
                "grouped mode: single-CTA 128-row tiles with direct stores, or split-N pair tiles (256 rows) staged through smem");
  static_assert(kEpi == 0 || (BLOCK_N == 256 && kTmaStore && kDiag == 0 && !kGrouped),
                "fused output quantisation: 256-wide tiles (one 128-column group per epilogue warp), TMA-store epilogue");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 128B-swizzled TMA/UMMA tiles need 1024-byte alignment.  The budget (kSmemOverhead) has no slack for a
  // misaligned base, so fail loudly rather than run off the end of the allocation.
  if ((ptx::smem_u32(smem_raw) & 1023u) != 0) __trap();
  const uint32_t smem_base = ptx::smem_u32(smem_raw);
  // the ring is sized on the host: `num_stages` stages of `stage_bytes` (A slot cut to the rows that
  // exist for decode shapes, so more weight bytes are in flight), at most kMaxStages
  const int nstages = p.num_stages;
  const uint32_t stage_bytes = p.stage_bytes;
  const uint32_t smem_store = smem_base + uint32_t(nstages) * stage_bytes;
  // kPipe == 3: TWO staging boxes per epilogue warp (one per 64-column half of its 128 columns), so the second TMA store
  // of a tile does not wait for the first one to have read the box — the per-tile store phase is what the MMA pipe idles on
  constexpr int kBoxes = (kPipe == 3) ? 2 : 1;
  const uint32_t bar_base = smem_store + (kTmaStore ? C::STORE_BYTES * kBoxes : 0);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + kAccBars + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 2 * kAccBars);
  // scale staging for the promotion warps (filled with cp.async so no long-latency LDG sits in their
  // in-order instruction stream): 2 x fp32 per warp for sb at bar_base + 512, 2 x fp32 per thread for sa after the KB of barriers
  const uint32_t ready_word = bar_base + 400u;   // gated mode: the producer's mask of arrived m-blocks, published for the promotion warps
  const uint32_t sb_slots = bar_base + 512u;
  const uint32_t sa_slots = bar_base + 1024u;
  uint8_t* smem_gen = smem_raw + (smem_base - ptx::smem_u32(smem_raw));   // generic pointer to smem_base

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (kCl == 2) ? ptx::cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_b);
    if (kTmaStore)
      for (int d = 0; d < p.num_dests; ++d) ptx::prefetch_tensormap(&tmap_c.m[d]);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < nstages; ++s) {
      ptx::mbar_init(full_bar(s), 1);                    // one arrive.expect_tx by the (leader's) producer
      ptx::mbar_init(empty_bar(s), kMcB ? 2 : 1);        // one tcgen05.commit (multicast-B: this CTA's and the peer's)
    }
    for (int s = 0; s < kAccBars; ++s) {
      ptx::mbar_init(tfull_bar(s), 1);                             // one tcgen05.commit
      ptx::mbar_init(tempty_bar(s), kCG * kNumEpiWarps / kHalves); // one arrive per epilogue warp (of that half) of every CTA
    }
    if constexpr (kWaitA) asm volatile("st.shared.u64 [%0], %1;" ::"r"(ready_word), "l"(0ull) : "memory");
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc<kCG>(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish<kCG>();
  }
  ptx::tc_fence_before();
  if (kCl == 2) ptx::cluster_sync(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  // m-blocks of 128*kCG rows; grouped: the m-tile table was built on the device from the expert offsets
  const int num_m_tiles = kGrouped ? __ldg(p.grp_num_m_tiles) : p.num_m_blocks;
  const int num_tiles = num_m_tiles * p.num_n_blocks;
  // grouped: tile -> (m-tile, n-block).  p.group_m == 0: the n-block fastest — the CTAs running together work on a few
  // m-tiles (a few experts): an m-tile's A panel is read once, and an expert's B panel is still in L2 when its next m-tile
  // comes by.  Otherwise (p.group_m != 0) the m-tile fastest: every wave touches every m-tile, and with more activations
  // than L2 holds they are re-read from HBM once per n-block.
  auto grouped_tile = [&](int tile, int& nb) {
    int mt;
    if (p.group_m == 0) { mt = tile / p.num_n_blocks; nb = tile - mt * p.num_n_blocks; }
    else { nb = tile / num_m_tiles; mt = tile - nb * num_m_tiles; }
    return __ldg(p.grp_tiles + mt);
  };
  (void)grouped_tile;
  const int num_workers = gridDim.x / kCl;
  const int worker = blockIdx.x / kCl;
  const int num_kb = p.num_k_blocks;

  // Register budget: 384 threads x 168 at launch; the 4 control warps give theirs back so the
  // promotion warps can hold 128 fp32 accumulators + two 32-column TMEM staging buffers.
  // (each role's code must be dominated by its own setmaxnreg, or ptxas budgets it for the minimum)
  if (warp == 0) {
    // =========================== TMA producer ===========================
    ptx::reg_dealloc<40>();
    {
      const uint64_t pol_a = ptx::policy_evict_normal();
      const uint64_t pol_b = ptx::policy_evict_normal();
      int stage = 0;
      uint32_t phase = 0;
      WorkIter it(p, worker, num_workers, num_tiles);
      int tile, kb0, kb1;
      while (it.next(tile, kb0, kb1)) {
        int mb, nb, m0, n0, n_tile;
        if constexpr (kGrouped) {
          const int4 ti = grouped_tile(tile, nb);
          m0 = ti.x + int(cta_rank) * BLOCK_M;                  // rows past the expert's range are loaded but never stored
          n_tile = ti.z * int(p.N) + nb * BLOCK_N;              // B is the experts' [N, K] matrices stacked
          n0 = n_tile + int(cta_rank) * C::LOAD_N;
          mb = 0; (void)mb;
        } else {
          if constexpr (kWaitA) {
            tile_coords_panels(tile, p, mb, nb);
            // Wait until this tile's m-block has arrived.  Fast path: one LDS + warp vote on the mask of arrived blocks;
            // refresh: ONE warp-wide round of system-scope acquire loads over all (<= 64) counters — such a load costs
            // ~1-2 us, so it must not run per tile once everything is there — then the proxy fence (the rows were written
            // by another GPU's generic-proxy stores, the loads below are async-proxy reads) and the mask is published for
            // the promotion warps.  Everything here is warp-CONVERGENT (all lanes, vote-controlled loops) and lives in the
            // producer: a lane-0 spin, a non-inlined helper, or a fourth warp role with its own tile loop each made
            // ptxas drop the uniform-register form of the MMA issuer's loop (7 R2UR per K block half: +12 % kernel time).
            if (p.num_m_blocks <= 64) {
              unsigned long long seen;
              asm volatile("ld.shared.u64 %0, [%1];" : "=l"(seen) : "r"(ready_word) : "memory");
              for (uint32_t spins = 0; !__all_sync(0xFFFFFFFFu, int((seen >> mb) & 1ull)); ++spins) {
                uint32_t v0 = p.a_ready_target - 1u, v1 = p.a_ready_target - 1u;
                if (lane < p.num_m_blocks)
                  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v0) : "l"(p.a_ready + lane) : "memory");
                if (lane + 32 < p.num_m_blocks)
                  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v1) : "l"(p.a_ready + lane + 32) : "memory");
                const uint32_t w0 = __ballot_sync(0xFFFFFFFFu, int32_t(v0 - p.a_ready_target) >= 0);
                const uint32_t w1 = __ballot_sync(0xFFFFFFFFu, int32_t(v1 - p.a_ready_target) >= 0);
                seen = uint64_t(w0) | (uint64_t(w1) << 32);
                asm volatile("fence.proxy.async;" ::: "memory");
                if (ptx::elect_one()) asm volatile("st.release.cta.shared.u64 [%0], %1;" ::"r"(ready_word), "l"(seen) : "memory");
                if (p.a_wait_polls != 0u && spins > p.a_wait_polls) __trap();   // caller-set bound on waiting for a peer (0: none)
              }
            } else {                                       // more than 64 m-blocks: this tile's own counter, every tile
              const uint32_t* cp = p.a_ready + mb;
              uint32_t v, spins = 0;
              do {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(cp) : "memory");
                if (++spins > p.a_wait_polls && p.a_wait_polls != 0u) __trap();
              } while (__any_sync(0xFFFFFFFFu, int32_t(v - p.a_ready_target) < 0));
              asm volatile("fence.proxy.async;" ::: "memory");
            }
          } else {
            tile_coords(tile, p, mb, nb);
          }
          m0 = mb * (BLOCK_M * kCl) + int(cta_rank) * BLOCK_M;
          n_tile = nb * BLOCK_N;
          n0 = n_tile + int(cta_rank) * (kMcB ? BLOCK_N / 2 : C::LOAD_N);
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa_addr = smem_base + stage * stage_bytes;
          const uint32_t sb_addr = sa_addr + p.a_slot_bytes;
          if (ptx::elect_one()) {
            if constexpr (kMcB) {
              // own A rows; this CTA's half of the B tile goes to BOTH CTAs (each barrier expects A + all of B)
              ptx::mbar_arrive_expect_tx(full_bar(stage), p.stage_tx_bytes);
              ptx::tma_load_2d(&tmap_a, full_bar(stage), sa_addr, kb * BLOCK_K, m0, pol_a);
              ptx::tma_load_2d_mcast(&tmap_b, full_bar(stage), sb_addr + cta_rank * uint32_t((BLOCK_N / 2) * BLOCK_K),
                                     kb * BLOCK_K, n0, uint16_t(3), pol_b);
            } else if constexpr (kCG == 1) {
              ptx::mbar_arrive_expect_tx(full_bar(stage), p.stage_tx_bytes);
              ptx::tma_load_2d(&tmap_a, full_bar(stage), sa_addr, kb * BLOCK_K, m0, pol_a);
              ptx::tma_load_2d(&tmap_b, full_bar(stage), sb_addr, kb * BLOCK_K, n0, pol_b);
            } else {
              // both CTAs' bytes are counted on the leader's barrier
              const uint32_t leader_full = full_bar(stage) & 0xFEFFFFFFu;
              if (is_leader) ptx::mbar_arrive_expect_tx(full_bar(stage), p.stage_tx_bytes);
              ptx::tma_load_2d_pair(&tmap_a, leader_full, sa_addr, kb * BLOCK_K, m0, pol_a);
              if constexpr (kSplitN) {
                // column half h of the pair's tile = [n_tile + h*128, +128): this CTA contributes 64 rows of each half
                const int nt = n_tile + int(cta_rank) * (BLOCK_N / 4);
                ptx::tma_load_2d_pair(&tmap_b, leader_full, sb_addr, kb * BLOCK_K, nt, pol_b);
                ptx::tma_load_2d_pair(&tmap_b, leader_full, sb_addr + (BLOCK_N / 4) * BLOCK_K, kb * BLOCK_K, nt + BLOCK_N / 2, pol_b);
              } else {
                ptx::tma_load_2d_pair(&tmap_b, leader_full, sb_addr, kb * BLOCK_K, n0, pol_b);
              }
            }
          }
          __syncwarp();
          if (++stage == nstages) { stage = 0; phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // =========================== MMA issuer (leader CTA only) ===========================
    ptx::reg_dealloc<40>();
    // The whole warp walks the loop convergently and one elected lane issues: descriptors, barrier
    // addresses and TMEM addresses then live in uniform registers (a lane-0-only branch made ptxas wrap
    // every tcgen05.mma in an ELECT/R2UR loop and the issuing thread became the bottleneck — profile r1b).
    if (is_leader || kMcB) {
      constexpr uint32_t idesc = make_idesc(BLOCK_M * kCG, BLOCK_N);
      const uint32_t tmem_base_u = __shfl_sync(0xFFFFFFFFu, tmem_base, 0);
      const uint64_t desc_hi = make_smem_desc(0);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      WorkIter it(p, worker, num_workers, num_tiles);
      int tile, kb0, kb1;
      while (it.next(tile, kb0, kb1)) {
        for (int kb = kb0; kb < kb1; ++kb) {
          if constexpr (kSplitN) {
            ptx::mbar_wait(full_bar(stage), phase);
            const uint32_t sa_addr = smem_base + stage * stage_bytes;
            const uint64_t da = desc_hi | uint64_t((sa_addr & 0x3FFFFu) >> 4);
            const uint64_t db = desc_hi | uint64_t(((sa_addr + p.a_slot_bytes) & 0x3FFFFu) >> 4);
            constexpr uint32_t idesc_h = make_idesc(BLOCK_M * kCG, BLOCK_N / 2);
            constexpr uint32_t kHalfBDesc = uint32_t((BLOCK_N / 2 / kCG) * BLOCK_K) >> 4;   // B rows of one half in this CTA
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              ptx::mbar_wait(tempty_bar(acc * 2 + h), acc_phase ^ 1);     // that half drained by its epilogue warps?
              if (kDiag == 4 && h == 0 && blockIdx.x == 0 && lane == 0 && kb < 64 && tile == worker) p.dbg[kb * 8 + 0] = clock64();
              ptx::tc_fence_after();
              const uint32_t d_tmem = tmem_base_u + uint32_t(acc * BLOCK_N + h * (BLOCK_N / 2));
              if (ptx::elect_one()) {
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                  ptx::mma_f8f6f4<kCG>(d_tmem, da + uint64_t(k * (UMMA_K >> 4)),
                                       db + uint64_t(h * kHalfBDesc + k * (UMMA_K >> 4)), idesc_h, k > 0 ? 1u : 0u);
                if (h == 1) ptx::tc_commit<kCG>(empty_bar(stage));
                ptx::tc_commit<kCG>(tfull_bar(acc * 2 + h));
              }
              __syncwarp();
            }
            if (kDiag == 4 && blockIdx.x == 0 && lane == 0 && kb < 64 && tile == worker) p.dbg[kb * 8 + 2] = clock64();
            if (++stage == nstages) { stage = 0; phase ^= 1; }
            if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
            continue;
          }
          // operands first (they arrive long before the accumulator frees up), descriptors built, and only then
          // the wait that sits on the critical ring: accumulator stage drained by every epilogue warp (of both CTAs)?
          ptx::mbar_wait(full_bar(stage), phase);
          const uint32_t sa_addr = smem_base + stage * stage_bytes;
          const uint64_t da = desc_hi | uint64_t((sa_addr & 0x3FFFFu) >> 4);
          const uint64_t db = desc_hi | uint64_t(((sa_addr + p.a_slot_bytes) & 0x3FFFFu) >> 4);
          if ((kDiag != 2 && kDiag != 3) || kb == kb0) ptx::mbar_wait(tempty_bar(acc), acc_phase ^ 1);
          if (kDiag >= 4 && blockIdx.x == 0 && lane == 0 && kb < 64 && tile == worker) p.dbg[kb * 8 + 0] = clock64();
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base_u + uint32_t(acc * BLOCK_N);
          const bool last_of_tile = (kDiag == 2 || kDiag == 3) && kb + 1 < kb1;
          if (ptx::elect_one()) {
            if constexpr (kDiag == 3) {   // diagnostic: unscaled GEMM issued as two N/2-wide halves (A read twice from smem)
              constexpr uint32_t idesc_h = make_idesc(BLOCK_M * kCG, BLOCK_N / 2);
#pragma unroll
              for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                  ptx::mma_f8f6f4<kCG>(d_tmem + uint32_t(h * (BLOCK_N / 2)), da + uint64_t(k * (UMMA_K >> 4)),
                                       db + uint64_t(h * ((BLOCK_N / 2 / kCG) * BLOCK_K >> 4) + k * (UMMA_K >> 4)), idesc_h,
                                       (k > 0 || kb > kb0) ? 1u : 0u);
            } else
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
              // +32 B along K inside the swizzle atom = +2 in the descriptor's 16-byte address units
              ptx::mma_f8f6f4<kCG>(d_tmem, da + uint64_t(k * (UMMA_K >> 4)), db + uint64_t(k * (UMMA_K >> 4)), idesc,
                                   (k > 0 || (kDiag == 2 && kb > kb0)) ? 1u : 0u);
            }
            // smem stage reusable once these MMAs have read it (multicast-B: the peer's producer writes here too)
            if constexpr (kMcB) ptx::tc_commit_mcast1(empty_bar(stage), uint16_t(3));
            else ptx::tc_commit<kCG>(empty_bar(stage));
            if (!last_of_tile) ptx::tc_commit<kCG>(tfull_bar(acc));   // partial for this K block complete
          }
          __syncwarp();
          if (kDiag >= 4 && blockIdx.x == 0 && lane == 0 && kb < 64 && tile == worker) p.dbg[kb * 8 + 1] = clock64();   // MMAs + commits issued
          if constexpr (kDiag == 10) {
            // diagnostic: fully serialised — wait for this block's own commit before going on: the period is then
            // issue + MMA execution + commit-to-observer latency
            ptx::mbar_wait(tfull_bar(acc), acc_phase);
            if (blockIdx.x == 0 && lane == 0 && kb < 64 && tile == worker) p.dbg[kb * 8 + 2] = clock64();
          } else
          if constexpr (kDiag >= 4) {
            // timeline diagnostic: when does the commit of the PREVIOUS K block's MMAs land, as seen by a spinning
            // thread?  (MMA(kb) is queued behind it, so this wait never delays the next issue.)
            if (kb > kb0) {
              const int pa = acc ^ 1;
              const uint32_t pp = (acc == 0) ? (acc_phase ^ 1) : acc_phase;
              ptx::mbar_wait(tfull_bar(pa), pp);
              if (blockIdx.x == 0 && lane == 0 && kb - 1 < 64 && tile == worker) p.dbg[(kb - 1) * 8 + 2] = clock64();
            }
          }
          if (++stage == nstages) { stage = 0; phase ^= 1; }
          if (last_of_tile) continue;
          if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp < 4) {
    ptx::reg_dealloc<40>();                  // warps 2 (TMEM owner) and 3 (spare) only wait for the end
  } else {
    // =========================== promotion + epilogue warps ===========================
    ptx::reg_alloc<232>();
    const int ew = warp - 4;
    const int quarter = warp & 3;            // TMEM lanes this warp may touch: 32*quarter .. +31
    const int half = ew >> 2;                // which half of the tile's columns
    const uint32_t lane_base = uint32_t(quarter * 32) << 16;
    const uint32_t tempty_leader = (kCG == 2) ? ptx::mapa(tempty_bar(0), 0) : tempty_bar(0);
    const uint32_t my_store0 = smem_store + ew * (32 * kStoreBoxCols * 2) * kBoxes;
    int acc = 0;
    uint32_t acc_phase = 0;
    WorkIter it(p, worker, num_workers, num_tiles);
    int tile, kb0, kb1;
    bool first_segment = true;
    while (it.next(tile, kb0, kb1)) {
      int mb, nb, m0;
      int row_limit = int(p.M);       // first row this tile must not write (grouped: the next expert's first row)
      bool row_ok;
      const float* sb_base = p.sb;
      if constexpr (kGrouped) {
        const int4 ti = grouped_tile(tile, nb);
        m0 = ti.x + int(cta_rank) * BLOCK_M;
        row_ok = int(cta_rank) * BLOCK_M + quarter * 32 + lane < ti.y;   // rows of this expert in the tile
        row_limit = ti.x + ti.y;
        sb_base += int64_t(ti.z) * p.grp_sb_expert_stride;
        mb = 0;
      } else {
        if constexpr (kWaitA) {
          // (no poll here: this warp's first read of the m-block's scales comes after the tile's first accumulator
          // hand-off, i.e. after the producer saw the counter — the writer made its rows visible before the counter)
          tile_coords_panels(tile, p, mb, nb);
        } else {
          tile_coords(tile, p, mb, nb);
        }
        m0 = mb * (BLOCK_M * kCl) + int(cta_rank) * BLOCK_M;
        row_ok = m0 + quarter * 32 + lane < p.M;
      }
      const int row = m0 + quarter * 32 + lane;
      const int col0 = nb * BLOCK_N + half * NC;          // first column this warp owns
      // scale pointers: sa walks kb for this row, sb walks kb for this warp's 128-column block
      const float* sa_ptr = p.sa + int64_t(row_ok ? row : 0) * p.sa_sm;
      const int sbc = min(col0 / 128, p.num_sb_cols - 1);
      const float* sb_ptr = sb_base + int64_t(sbc) * p.sb_sn;
      // NC <= 128 and col0 % NC == 0, so the warp's columns never straddle a 128-wide scale block
      float accum[NC];
#pragma unroll
      for (int i = 0; i < NC; ++i) accum[i] = 0.0f;
      // Scale prefetch, two K blocks ahead, kept as raw loads: the product is formed only at the top
      // of the iteration that uses it, so no instruction ever waits on an in-flight global load
      // (an in-order warp that multiplies right after the load stalls ~1 us per K block — measured).
      const int64_t sa_sk = p.sa_sk, sb_sk = p.sb_sk;
      // (gated mode: the scales were written by other GPUs during this launch, so not through the non-coherent path; a
      // plain L1-cached load is safe — L1 starts every launch empty and nothing of an m-block is read before its counter
      // has been seen — and needed: the next group's scales are requested ~300 clk before their use, an L1 hit makes
      // that, an L2 round trip does not: ld.cg cost 7 % of the kernel)
      auto ld_sa = [&](int kb) {
        return (row_ok && kb < kb1) ? (kWaitA ? *(sa_ptr + int64_t(kb) * sa_sk) : __ldg(sa_ptr + int64_t(kb) * sa_sk)) : 0.0f;
      };
      auto ld_sa4 = [&](const float4* ptr) {
        if constexpr (kWaitA) return *ptr;      // (a plain load the compiler may schedule, like the __ldg of the other form)
        else return __ldg(ptr);
      };
      auto ld_sb = [&](int kb) { return kb < kb1 ? __ldg(sb_ptr + int64_t(kb) * sb_sk) : 0.0f; };
      float sa0 = ld_sa(kb0), sb0 = ld_sb(kb0), sa1 = ld_sa(kb0 + 1), sb1 = ld_sb(kb0 + 1);
      constexpr int NCH = NC / 32;           // 32-column TMEM chunks per K block for this warp
      const int hb = kSplitN ? half : 0;     // which hand-off barrier of the stage this warp uses
      auto release_stage = [&]() {
        // every tcgen05.ld of this warp on the stage has completed: hand it back to the MMA issuer
        // (tcgen05.wait::ld is warp-aligned: every lane's loads are complete here, no extra warp sync needed)
        ptx::tc_fence_before();
        if (lane == 0) {
          if constexpr (kCG == 2) ptx::mbar_arrive_cluster(tempty_leader + 8u * (acc * kHalves + hb));
          else ptx::mbar_arrive(tempty_bar(acc * kHalves + hb));
        }
        if (++acc == kAccStages) { acc = 0; acc_phase ^= 1; }
      };
      auto stage_addr = [&]() { return tmem_base + lane_base + uint32_t(acc * BLOCK_N + half * NC); };
      uint32_t v[2][32];
      if constexpr (kDiag == 9) {   // diagnostic: immediate hand-off, then the promotion's FMA volume on stale registers
        ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
        ptx::tc_fence_after();
        ptx::tmem_ld_32x32(stage_addr(), v[0]);
        ptx::tmem_ld_32x32(stage_addr() + 32u, v[1]);
        ptx::tmem_ld_wait(v[0]);
        ptx::tmem_ld_wait(v[1]);
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          ptx::tc_fence_after();
          release_stage();
          const float s = sa0 * sb0 + float(kb);
#pragma unroll
          for (int c = 0; c < NCH; ++c)
#pragma unroll
            for (int i = 0; i < 32; i += 2)
              ptx::fma2(accum[c * 32 + i], accum[c * 32 + i + 1], v[c & 1][i], v[c & 1][i + 1], s);
        }
      } else if constexpr (kDiag == 1 || kDiag == 5 || kDiag == 10) {            // diagnostic: TMA + MMA pipeline with the per-block hand-off, no arithmetic
        for (int kb = kb0; kb < kb1; ++kb) {
          ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          ptx::tc_fence_after();
          accum[0] += sa0 * sb0 + float(stage_addr());
          release_stage();
        }
      } else if constexpr (kDiag == 2 || kDiag == 3) {     // diagnostic: unscaled GEMM, one hand-off per tile
        ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
        ptx::tc_fence_after();
        const uint32_t taddr = stage_addr();
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          ptx::tmem_ld_32x32(taddr + uint32_t(c * 32), v[0]);
          ptx::tmem_ld_wait(v[0]);
#pragma unroll
          for (int i = 0; i < 32; ++i) accum[c * 32 + i] = __uint_as_float(v[0][i]) * (sa0 * sb0);
        }
        release_stage();
      } else if constexpr (NCH % 2 == 0) {
        // Software pipeline across K blocks: chunk c+1 is in flight while chunk c is folded in, the
        // stage is released as soon as its last chunk is in registers (before that chunk's FMAs),
        // and the first chunk of the NEXT K block is requested before the last FMAs of this one.
        int dbg_kb = kb0;
        // Scales travel global -> smem with cp.async (4 bytes per thread for sa, per warp for sb), two K blocks
        // ahead, and are read back with LDS: a plain LDG here shares the 6 scoreboard slots with the tcgen05.ld
        // stream and its L2 latency then stalls whatever instruction reuses the slot (measured 250-400 clk per block).
        uint32_t taddr = 0;
        const uint32_t my_sa = sa_slots + uint32_t(threadIdx.x - 128) * 8u;
        const uint32_t my_sb = sb_slots + uint32_t(ew) * 8u;
        // running global pointers (bumped by one K block per prefetch: nothing to rematerialise) and a slot toggle
        const float* sa_next = sa_ptr + int64_t(kb0) * sa_sk;
        const float* sb_next = sb_ptr + int64_t(kb0) * sb_sk;
        int kb_next = kb0;
        uint32_t fill_off = 0, read_off = 0;             // byte offset (0 or 4) of the slot to fill / to read next
        auto prefetch_scales = [&]() {
          if (kb_next < kb1) {
            if (row_ok) ptx::cp_async4(my_sa + fill_off, sa_next);
            if (lane == 0) ptx::cp_async4(my_sb + fill_off, sb_next);
          }
          ptx::cp_async_commit();
          sa_next += sa_sk; sb_next += sb_sk; ++kb_next; fill_off ^= 4u;
        };
        auto staged_scale = [&]() {
          ptx::cp_async_wait<1>();                       // the group of this K block has landed (one more may be in flight)
          if (kDiag == 4 && blockIdx.x == 0 && ew == 0 && lane == 0 && dbg_kb < 64 && tile == worker) p.dbg[dbg_kb * 8 + 2] = clock64();
          __syncwarp();                                  // lane 0's sb copy visible to the warp
          const float s = ptx::lds_f32(my_sa + read_off) * ptx::lds_f32(my_sb + read_off);
          read_off ^= 4u;
          __syncwarp();                                  // everyone has read the slot before it is refilled
          prefetch_scales();
          return s;
        };
        float4 a_nx, b_nx;                               // vector path: the next group's raw scales
        const float4* sa4n = nullptr;
        const float4* sb4n = nullptr;
        // kPipe == 1: chunk loads go out in PAIRS (both staging buffers), so a K block costs two tcgen05.ld round trips
        // instead of four and only two chunks of FMAs (not three) stand between the block's commit and the hand-back
        // of its accumulator stage.  Entry state: chunks 0 and 1 of this block are already in flight.
        auto fma_chunk = [&](int c, const uint32_t (&buf)[32], float s) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) ptx::fma2(accum[c * 32 + i], accum[c * 32 + i + 1], buf[i], buf[i + 1], s);
        };
        auto promote2 = [&](const float s, const bool more, const bool fetch = false) {
          static_assert(kPipe != 1 || NCH == 4, "paired chunk loads: 128 columns per epilogue warp");
          static_assert(kPipe != 2 || NCH == 2, "whole-share staging: 64 columns per epilogue warp");
          ptx::tmem_ld_wait(v[0]);
          ptx::tmem_ld_wait(v[1]);
          fma_chunk(0, v[0], s);
          ptx::tmem_ld_32x32(taddr + 64u, v[0]);
          fma_chunk(1, v[1], s);
          ptx::tmem_ld_32x32(taddr + 96u, v[1]);
          ptx::tmem_ld_wait(v[0]);
          ptx::tmem_ld_wait(v[1]);
          release_stage();
          bool ready = false;
          if (more) {
            taddr = stage_addr();
            ready = ptx::mbar_try_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          }
          if (fetch) {
            a_nx = ld_sa4(sa4n);
            b_nx = __ldg(sb4n);
          }
          fma_chunk(2, v[0], s);
          if (more) {
            if (!ready) ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
            ptx::tc_fence_after();
            ptx::tmem_ld_32x32(taddr, v[0]);
          }
          fma_chunk(3, v[1], s);
          if (more) ptx::tmem_ld_32x32(taddr + 32u, v[1]);
        };
        // kPipe == 2 (128-wide tiles: 64 columns per epilogue warp = both staging buffers): the WHOLE share of the K block
        // is parked in registers, the accumulator stage goes back to the MMA issuer at once and all FMAs come after —
        // the order the diagnostics say costs nothing (hand-back first, FMAs after).  Entry state: both chunks in flight.
        auto promote3 = [&](const float s, const bool more, const bool fetch = false) {
          ptx::tmem_ld_wait(v[0]);
          ptx::tmem_ld_wait(v[1]);
          release_stage();
          bool ready = false;
          if (more) {
            taddr = stage_addr();
            ready = ptx::mbar_try_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          }
          if (fetch) {
            a_nx = ld_sa4(sa4n);
            b_nx = __ldg(sb4n);
          }
          fma_chunk(0, v[0], s);
          if (more) {
            if (!ready) ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
            ptx::tc_fence_after();
            ptx::tmem_ld_32x32(taddr, v[0]);
          }
          fma_chunk(1, v[1], s);
          if (more) ptx::tmem_ld_32x32(taddr + 32u, v[1]);
        };
        auto promote = [&](const float s, const bool more, const bool fetch = false) {
          if constexpr (kPipe == 1) { promote2(s, more, fetch); return; }
          if constexpr (kPipe == 2) { promote3(s, more, fetch); return; }
          const bool stamp = (kDiag == 4 || kDiag >= 6) && blockIdx.x == 0 && ew == 0 && lane == 0 && dbg_kb < 64 && tile == worker;
#pragma unroll
          for (int c = 0; c < NCH; ++c) {
            if constexpr (kDiag != 7 && kDiag != 8) ptx::tmem_ld_wait(v[c & 1]);
            if (stamp && c == 0) p.dbg[dbg_kb * 8 + 3] = clock64();       // first chunk of this K block in registers
            if (c + 1 < NCH) {
              if constexpr (kDiag != 7 && kDiag != 8) ptx::tmem_ld_32x32(taddr + uint32_t((c + 1) * 32), v[(c + 1) & 1]);
            } else {
              if (stamp) p.dbg[dbg_kb * 8 + 4] = clock64();                 // last chunk in registers: stage released
              release_stage();
              if (stamp) p.dbg[dbg_kb * 8 + 5] = clock64();                 // after fence + arrive
              // Probe the next accumulator's barrier, fold in the first half of this chunk while the probe is in
              // flight (its ~100-clk round trip is otherwise dead time in this in-order warp), then request the next
              // block's first chunk and finish the second half under that load.
              bool ready = false;
              if (more) {
                taddr = stage_addr();
                ready = ptx::mbar_try_wait(tfull_bar(acc * kHalves + hb), acc_phase);
              }
              // the next group's scales are requested here, in the shadow of the hand-off wait, not at the group's
              // loop top where the issue sits between this warp's first tcgen05.ld and its FMAs
              if (fetch) {
                a_nx = ld_sa4(sa4n);
                b_nx = __ldg(sb4n);
              }
              if (kDiag != 6 && !(kDiag == 8 && (warp & 3) == 1))
#pragma unroll
              for (int i = 0; i < 16; i += 2)
                ptx::fma2(accum[c * 32 + i], accum[c * 32 + i + 1], v[c & 1][i], v[c & 1][i + 1], s);
              if (more) {
                if (!ready) ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
                ptx::tc_fence_after();
                if constexpr (kDiag != 7 && kDiag != 8) ptx::tmem_ld_32x32(taddr, v[0]);
                if (stamp) p.dbg[dbg_kb * 8 + 7] = clock64();               // next block's first tcgen05.ld issued
              }
              if (kDiag != 6 && !(kDiag == 8 && (warp & 3) == 1))
#pragma unroll
              for (int i = 16; i < 32; i += 2)
                ptx::fma2(accum[c * 32 + i], accum[c * 32 + i + 1], v[c & 1][i], v[c & 1][i + 1], s);
              continue;
            }
            if (kDiag != 6 && !(kDiag == 8 && (warp & 3) == 1))
#pragma unroll
            for (int i = 0; i < 32; i += 2)
              ptx::fma2(accum[c * 32 + i], accum[c * 32 + i + 1], v[c & 1][i], v[c & 1][i + 1], s);
            else accum[c] += s;
          }
          if (stamp) p.dbg[dbg_kb * 8 + 6] = clock64();                     // FMAs of this K block done
          ++dbg_kb;
        };
        // two K blocks per trip so the scale registers rotate by name (no register moves that would
        // touch a load still in flight): each pair is reloaded right after its use, two blocks ahead
        if (p.scale_vec) {
          // Scales contiguous along K blocks (row-major sa as the quant op writes by default; sb as the
          // weight-scale view gives it): one 16-byte load per FOUR K blocks per thread, fetched a whole group
          // ahead into registers — no per-block load, wait or smem round trip in the promotion loop.
          const float4* sa4 = reinterpret_cast<const float4*>(sa_ptr);
          const float4* sb4 = reinterpret_cast<const float4*>(sb_ptr);
          int g = kb0 >> 2;
          const int g_end = (kb1 + 3) >> 2;
          // (rows >= M read row 0's scales: in bounds, and those accumulator rows are never stored)
          // running pointers, bumped by 16 bytes per group: anything derived from g is rematerialised by ptxas at
          // every group boundary under this register pressure (S2R + LDCU + 64-bit IMAD chain, ~200 clk measured)
          sa4n = sa4 + g;
          sb4n = sb4 + g;
          // gated mode: the scales may be requested ahead of the tile's first hand-off only if the producer has already
          // published the m-block as arrived; otherwise after the hand-off (which the producer's own wait precedes)
          bool early = true;
          if constexpr (kWaitA) {
            unsigned long long seen;
            asm volatile("ld.acquire.cta.shared.u64 %0, [%1];" : "=l"(seen) : "r"(ready_word) : "memory");
            early = p.num_m_blocks <= 64 && ((seen >> mb) & 1ull);
          }
          if (early) a_nx = ld_sa4(sa4n);
          b_nx = __ldg(sb4n);
          ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          if (!early) a_nx = ld_sa4(sa4n);
          ptx::tc_fence_after();
          taddr = stage_addr();
          ptx::tmem_ld_32x32(taddr, v[0]);
          if constexpr (kPipe == 1 || kPipe == 2) ptx::tmem_ld_32x32(taddr + 32u, v[1]);
          if constexpr (kDiag == 7 || kDiag == 8) {     // diagnostic: FMAs on two chunks fetched once, no tcgen05.ld in the loop
            ptx::tmem_ld_32x32(taddr + 32u, v[1]);
            ptx::tmem_ld_wait(v[0]);
            ptx::tmem_ld_wait(v[1]);
          }
          for (; g < g_end; ++g) {
            const float s0 = a_nx.x * b_nx.x, s1 = a_nx.y * b_nx.y, s2 = a_nx.z * b_nx.z, s3 = a_nx.w * b_nx.w;
            ++sa4n; ++sb4n;
            const int kb = g << 2;
            if (kb >= kb0 && kb < kb1) promote(s0, kb + 1 < kb1);
            if (kb + 1 >= kb0 && kb + 1 < kb1) promote(s1, kb + 2 < kb1);
            if (kb + 2 >= kb0 && kb + 2 < kb1) promote(s2, kb + 3 < kb1);
            // a following group exists exactly when block kb+4 is in range, and then block kb+3 is too
            if (kb + 3 >= kb0 && kb + 3 < kb1) promote(s3, kb + 4 < kb1, kb + 4 < kb1);
          }
        } else {
          prefetch_scales();
          prefetch_scales();
          ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          ptx::tc_fence_after();
          taddr = stage_addr();
          ptx::tmem_ld_32x32(taddr, v[0]);
          if constexpr (kPipe == 1 || kPipe == 2) ptx::tmem_ld_32x32(taddr + 32u, v[1]);
          for (int kb = kb0; kb < kb1; ++kb) promote(staged_scale(), kb + 1 < kb1);
          ptx::cp_async_wait<0>();
        }
      } else {
        for (int kb = kb0; kb < kb1; ++kb) {
          const float s = sa0 * sb0;
          sa0 = sa1; sb0 = sb1;
          sa1 = ld_sa(kb + 2); sb1 = ld_sb(kb + 2);
          ptx::mbar_wait(tfull_bar(acc * kHalves + hb), acc_phase);
          ptx::tc_fence_after();
          const uint32_t taddr = stage_addr();
          ptx::tmem_ld_32x32(taddr, v[0]);
          ptx::tmem_ld_wait(v[0]);
#pragma unroll
          for (int c = 0; c < NCH; ++c) {
            if (c + 1 < NCH) ptx::tmem_ld_32x32(taddr + uint32_t((c + 1) * 32), v[(c + 1) & 1]);
            else release_stage();
#pragma unroll
            for (int i = 0; i < 32; i += 2)
              ptx::fma2(accum[c * 32 + i], accum[c * 32 + i + 1], v[c & 1][i], v[c & 1][i + 1], s);
            if (c + 1 < NCH) ptx::tmem_ld_wait(v[(c + 1) & 1]);
          }
        }
      }
      // ---------------- stream-K: a tile shared with other workers is completed by the last arriver ----------------
      bool write_out = true;
      if (p.stream_k && (kb0 != 0 || kb1 != num_kb)) {
        const int64_t total = int64_t(num_tiles) * num_kb;
        auto range_begin = [&](int w) { return total * w / num_workers; };
        auto owner = [&](int64_t u) { return int(((u + 1) * num_workers + total - 1) / total) - 1; };
        const int c_first = owner(int64_t(tile) * num_kb), c_last = owner(int64_t(tile + 1) * num_kb - 1);
        auto slot_of = [&](int c) { return 2 * c + ((int(range_begin(c) / num_kb) == tile) ? 0 : 1); };
        constexpr int kWarpFloats = NC * 32;                         // one warp's 32 x NC partial
        auto part_ptr = [&](int c) {
          return p.ws_partials + (int64_t(slot_of(c)) * (kCl * kNumEpiWarps) + int(cta_rank) * kNumEpiWarps + ew) * kWarpFloats;
        };
        float4* mine = reinterpret_cast<float4*>(part_ptr(worker));
        int32_t* cnt = p.ws_counters + int64_t(tile) * (kCl * kNumEpiWarps) + int(cta_rank) * kNumEpiWarps + ew;
        const int n_others = c_last - c_first;
        // PEEK at the counter first where that can succeed: the worker that holds a tile's first K blocks (kb0 == 0) reaches
        // it at the very end of its range, when everybody else has long arrived — it then knows it is the finisher, skips
        // the store + fence + atomic round trips of its own partial (they sit fully exposed at the end of the kernel) and,
        // being the first contributor in CTA order, simply adds the others to its registers in order (same bits as
        // 0 + own + others).  Not for a worker's first piece of work: contributors that start together (N=2112: four per
        // tile, one segment each) would only pay for the peek.  N=24576 decode: -2..-8 % GPU-side (tools/session30.sh).
        bool peeked = false;
        if (!first_segment && kb0 == 0) {
          int seen = 0;
          if (lane == 0) seen = *reinterpret_cast<volatile int32_t*>(cnt);
          peeked = __shfl_sync(0xFFFFFFFFu, seen, 0) == n_others;
        }
        bool finisher = peeked;
        if (!peeked) {
          if (row_ok) {
#pragma unroll
            for (int j = 0; j < NC / 4; ++j)
              __stcg(mine + j * 32 + lane, make_float4(accum[4 * j], accum[4 * j + 1], accum[4 * j + 2], accum[4 * j + 3]));
          }
          // release pattern: every lane's partial is ordered before lane 0 by the warp barrier, lane 0's (cumulative)
          // fence orders them before its counter increment
          __syncwarp();
          int arrived = 0;
          if (lane == 0) {
            // (an acq_rel fence, not __threadfence(): that is the sequentially consistent fence, which measured ~6 us
            // under a full TMA ring in the decode kernels — release / acquire is all this hand-over needs)
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            arrived = atomicAdd(cnt, 1);
          }
          arrived = __shfl_sync(0xFFFFFFFFu, arrived, 0);
          finisher = arrived == n_others;                            // everyone else's partial is already visible
        }
        write_out = finisher;
        if (write_out) {
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          if (lane == 0) *cnt = 0;                                   // leave the counters zeroed for the next launch
          // fixed order (bit-reproducible): 0 + partial(c_first) + partial(c_first + 1) + ...; after a successful peek
          // this worker IS c_first and its partial is still in registers
#pragma unroll
          for (int i = 0; i < NC; ++i) accum[i] = peeked ? 0.0f + accum[i] : 0.0f;
          for (int c = peeked ? c_first + 1 : c_first; c <= c_last; ++c) {
            const float4* src = reinterpret_cast<const float4*>(part_ptr(c));
            if (row_ok) {
#pragma unroll
              for (int j = 0; j < NC / 4; ++j) {
                const float4 t = __ldcg(src + j * 32 + lane);
                accum[4 * j] += t.x; accum[4 * j + 1] += t.y; accum[4 * j + 2] += t.z; accum[4 * j + 3] += t.w;
              }
            }
          }
        }
      }
      first_segment = false;
      // ---------------- write the finished 32 x NC sub-tile ----------------
      if (!write_out) continue;
      if constexpr (kEpi != 0) {
        // ---- fused output quantisation: this thread holds one whole 1x128 group (row `row`, columns col0..col0+127) ----
        // Bit-identical to writing C in OutT and running the quant op on it: round to OutT first, then the op's own
        // arithmetic (amax, RN(amax/448), the exhaustively verified RN(y/s) scheme of quant_kernel.cuh, sign copied).
#pragma unroll
        for (int i = 0; i < NC; ++i) accum[i] = round_to<OutT>(accum[i]);
        bool emit = col0 < p.N;                                   // warp-uniform
        int out_col = col0;
        if constexpr (kEpi == 2) {
          // gate | up: the up-half warp (ew + 4, same rows) parks its OutT-rounded values in the two staging boxes of
          // the pair; the gate-half warp folds them in as silu(g) * u and quantises.  Two named-barrier phases per
          // tile and pair: "boxes free" (the gate warp's previous TMA store has read its box) and "up values written".
          const uint32_t box_lo = smem_store + uint32_t(quarter) * (32 * kStoreBoxCols * 2);
          const uint32_t box_hi = box_lo + 4u * (32 * kStoreBoxCols * 2);
          const int bar_id = 1 + quarter;
          if (half == 1) {
            asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
#pragma unroll
            for (int j = 0; j < 16; ++j) {                       // 16 x 16 B = 128 OutT values of this lane's row
              const float* a = accum + j * 8;
              const uint32_t w0 = pack2<OutT>(a[0], a[1]), w1 = pack2<OutT>(a[2], a[3]);
              const uint32_t w2 = pack2<OutT>(a[4], a[5]), w3 = pack2<OutT>(a[6], a[7]);
              const uint32_t dst = (j < 8 ? box_lo : box_hi) + uint32_t(lane * 128 + (((j & 7) ^ (lane & 7)) << 4));
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
            }
            asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
            emit = false;
          } else {
            if (lane == 0) ptx::tma_store_wait_read<0>();
            __syncwarp();
            asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
            asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              uint32_t w[4];
              const uint32_t src = (j < 8 ? box_lo : box_hi) + uint32_t(lane * 128 + (((j & 7) ^ (lane & 7)) << 4));
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(src));
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const float2 u = unpack2<OutT>(w[t]);
                accum[j * 8 + 2 * t] = silu_mul<OutT>(accum[j * 8 + 2 * t], u.x);
                accum[j * 8 + 2 * t + 1] = silu_mul<OutT>(accum[j * 8 + 2 * t + 1], u.y);
              }
            }
            __syncwarp();                                        // every lane has read the boxes before box_lo is reused below
            out_col = nb * (BLOCK_N / 2);
          }
        }
        if (emit) {
          float amax = 0.0f;
#pragma unroll
          for (int i = 0; i < NC; ++i) amax = fmaxf(amax, fabsf(accum[i]));
          amax = fmaxf(amax, p.q_eps);
          // RN(amax / 448) exactly as the quant op forms it (same expression, so overflowed f16 groups agree too)
          constexpr float kRcp448 = 1.0f / 448.0f;
          const float q0s = amax * kRcp448;
          const float scale = __fmaf_rn(__fmaf_rn(-q0s, 448.0f, amax), kRcp448, q0s);
          if (row_ok) p.q_scales[int64_t(row) * p.qs_sm + int64_t(out_col >> 7) * p.qs_sg] = scale;
          const float r = __frcp_rn(scale), ns = -scale;
          if constexpr (kEpi == 1) {
            if (lane == 0) ptx::tma_store_wait_read<0>();        // the previous tile's store has read this warp's box
            __syncwarp();
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {                          // 8 x 16 B = 128 codes of this lane's row
            uint32_t w[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const float* a = accum + j * 16 + t * 4;
              const float2 c0 = div_pair(make_float2(a[0], a[1]), r, ns), c1 = div_pair(make_float2(a[2], a[3]), r, ns);
              const uint32_t lo = __nv_cvt_float2_to_fp8x2(c0, __NV_SATFINITE, __NV_E4M3);
              const uint32_t hi = __nv_cvt_float2_to_fp8x2(c1, __NV_SATFINITE, __NV_E4M3);
              const uint32_t mag = __byte_perm(lo, hi, 0x5410);
              const uint32_t s01 = __byte_perm(__float_as_uint(a[0]), __float_as_uint(a[1]), 0x0073);
              const uint32_t s23 = __byte_perm(__float_as_uint(a[2]), __float_as_uint(a[3]), 0x0073);
              const uint32_t sgn = __byte_perm(s01, s23, 0x5410);
              asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(w[t]) : "r"(sgn), "r"(mag), "r"(0x80808080u));   // sign ? y : code
            }
            const uint32_t dst = my_store0 + uint32_t(lane * 128 + ((j ^ (lane & 7)) << 4));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
          }
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_2d(&tmap_c.m[0], my_store0, out_col, m0 + quarter * 32);   // 32 rows x 128 codes, clipped at M / N
            ptx::tma_store_commit();
          }
        }
      } else
      if constexpr (kTmaStore) {
#pragma unroll
        for (int cb = 0; cb < NC / kStoreBoxCols; ++cb) {
          if (col0 + cb * kStoreBoxCols < p.N) {              // warp-uniform
            // the previous TMA store must have finished reading this warp's staging box
            const uint32_t my_store = my_store0 + uint32_t(kBoxes == 2 ? cb * (32 * kStoreBoxCols * 2) : 0);
            if (kBoxes == 1 || cb == 0) {
              if (lane == 0) ptx::tma_store_wait_read<0>();
              __syncwarp();
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {                     // 8 x 16 B = 64 bf16 of this lane's row
              const float* a = accum + cb * kStoreBoxCols + j * 8;
              const uint32_t w0 = pack2<OutT>(a[0], a[1]), w1 = pack2<OutT>(a[2], a[3]);
              const uint32_t w2 = pack2<OutT>(a[4], a[5]), w3 = pack2<OutT>(a[6], a[7]);
              const uint32_t dst = my_store + uint32_t(lane * 128 + ((j ^ (lane & 7)) << 4));
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
            }
            if (kGrouped || p.mc_ptr != nullptr) {
              // NVLS multicast form of the fused all-gather: one multimem.st per 16 bytes lands in EVERY GPU's
              // buffer (the NVSwitch replicates), so a rank sends its panel once instead of once per peer.
              // Grouped (MoE) form: plain 16-byte stores, rows cut at the expert's last row (a TMA box would run into
              // the next expert's rows).  Both read the staged box back row-contiguously: 8 lanes cover one 128-byte
              // row segment.
              __syncwarp();
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int r = i * 4 + (lane >> 3);
                const int ch = lane & 7;
                uint32_t w0, w1, w2, w3;
                const uint32_t src = my_store + uint32_t(r * 128 + ((ch ^ (r & 7)) << 4));
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "r"(src));
                const int grow = m0 + quarter * 32 + r;
                const int gcol = col0 + cb * kStoreBoxCols + ch * 8;
                if (grow < row_limit && gcol + 8 <= p.N) {
                  if constexpr (kGrouped) {
                    *reinterpret_cast<uint4*>(reinterpret_cast<OutT*>(p.C) + int64_t(grow) * p.ldc + gcol) = make_uint4(w0, w1, w2, w3);
                  } else {
                    OutT* dst = reinterpret_cast<OutT*>(p.mc_ptr) + int64_t(grow) * p.ldc + gcol;
                    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                                 ::"l"(dst), "f"(__uint_as_float(w0)), "f"(__uint_as_float(w1)), "f"(__uint_as_float(w2)),
                                   "f"(__uint_as_float(w3)) : "memory");
                  }
                }
              }
              __syncwarp();
            } else {
              ptx::fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                // one store per destination: this GPU's output and, for the fused all-gather, the same panel
                // position in every peer's [M, N] buffer (peer-mapped pointers, written over NVLink)
                for (int d = 0; d < p.num_dests; ++d)
                  ptx::tma_store_2d(&tmap_c.m[d], my_store, col0 + cb * kStoreBoxCols, m0 + quarter * 32);
                ptx::tma_store_commit();
              }
            }
          }
        }
      } else {
        OutT* crow = reinterpret_cast<OutT*>(p.C) + int64_t(row) * p.ldc;
        if (row_ok) {
          const bool vec_ok = (p.ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
#pragma unroll
          for (int j = 0; j < NC / 8; ++j) {
            const int col = col0 + j * 8;
            if (vec_ok && col + 8 <= p.N) {
              const float* a = accum + j * 8;
              uint4 w;
              w.x = pack2<OutT>(a[0], a[1]); w.y = pack2<OutT>(a[2], a[3]);
              w.z = pack2<OutT>(a[4], a[5]); w.w = pack2<OutT>(a[6], a[7]);
              *reinterpret_cast<uint4*>(crow + col) = w;
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i)
                if (col + i < p.N) {
                  const uint32_t w = pack2<OutT>(accum[j * 8 + i], 0.0f);
                  crow[col + i] = *reinterpret_cast<const OutT*>(&w);
                }
            }
          }
        }
      }
    }
    if (kTmaStore && lane == 0) ptx::tma_store_wait<0>();     // global writes complete before exit
    __syncwarp();
  }

  // =========================== teardown ===========================
  ptx::tc_fence_before();
  if (kCl == 2) ptx::cluster_sync(); else __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<kCG>(tmem_base, C::TMEM_COLS);
}

// Re-lays two strided [rows, cols] fp32 scale matrices out as dst[r * cols + c] (32x32 smem tiles, coalesced on both
// sides), one launch for both.  Used to turn M-major activation scales / [G, NB] weight scales into the
// K-block-contiguous form the promotion warps read with one 16-byte load per four K blocks; 3.7 MB at M=16384 —
// microseconds next to the GEMM.
struct RelayoutJob { const float* src; float* dst; int64_t s_r, s_c; int rows, cols, tiles_r; };
__global__ void __launch_bounds__(256) relayout_scales_kernel(RelayoutJob ja, RelayoutJob jb, int tiles_a) {
  __shared__ float tile[32][33];
  const bool second = int(blockIdx.x) >= tiles_a;
  const RelayoutJob& j = second ? jb : ja;
  const int t = second ? int(blockIdx.x) - tiles_a : int(blockIdx.x);
  const int r0 = (t % j.tiles_r) * 32, c0 = (t / j.tiles_r) * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
  const bool src_row_fast = j.s_r <= j.s_c;                          // which source index is (closer to) contiguous
  for (int i = ty; i < 32; i += 8) {
    const int r = src_row_fast ? r0 + tx : r0 + i, c = src_row_fast ? c0 + i : c0 + tx;
    if (r < j.rows && c < j.cols) tile[r - r0][c - c0] = j.src[int64_t(r) * j.s_r + int64_t(c) * j.s_c];
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int r = r0 + i, c = c0 + tx;
    if (r < j.rows && c < j.cols) j.dst[int64_t(r) * j.cols + c] = tile[i][tx];
  }
}

// Grouped (MoE) GEMM: the m-tile table.  Expert e owns rows [offsets[e], offsets[e] + sizes[e][0]) of the stacked A;
// it contributes ceil(m_e / tile_rows) entries {first row, valid rows, expert, 0} (tile_rows = 128, or 256 for pair tiles).  One block, one thread per expert,
// an inclusive scan of the tile counts in shared memory; everything stays on the device (no host sync, capturable).
__global__ void __launch_bounds__(1024)
build_group_tiles_kernel(const int32_t* __restrict__ offsets, const int32_t* __restrict__ sizes, int num_experts,
                         int4* __restrict__ table, int32_t* __restrict__ count, int max_tiles, int tile_rows) {
  __shared__ int scan[1024];
  int base = 0;                                           // tiles of the experts handled in earlier rounds
  for (int e0 = 0; e0 < num_experts; e0 += 1024) {
    const int e = e0 + int(threadIdx.x);
    const int m = e < num_experts ? max(sizes[e * 3], 0) : 0;
    const int nt = (m + tile_rows - 1) / tile_rows;
    scan[threadIdx.x] = nt;
    __syncthreads();
    for (int d = 1; d < 1024; d <<= 1) {
      const int v = int(threadIdx.x) >= d ? scan[threadIdx.x - d] : 0;
      __syncthreads();
      scan[threadIdx.x] += v;
      __syncthreads();
    }
    const int first = base + scan[threadIdx.x] - nt;
    if (e < num_experts) {
      const int off = offsets[e];
      for (int t = 0; t < nt; ++t)
        if (first + t < max_tiles)
          table[first + t] = make_int4(off + t * tile_rows, min(tile_rows, m - t * tile_rows), e, 0);
    }
    base += scan[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = min(base, max_tiles);
}

// --------------------------------------------------------------------------------------------
// host side
// --------------------------------------------------------------------------------------------
struct Variant {
  const char* name;
  int block_n, stages, cg;   // cg: CTAs that share one 128*cg-row tile slot (cluster size)
  bool tma_store;
  bool mcb;                  // cg == 2 as a cluster of two independent cta_group::1 tiles with multicast B
  bool product;              // false: laboratory variant, compiled only with -DSGLK_DIAG (libsglk_diag.so, tools/ only)
};
// index == variant id of the C ABI.  Ids are stable; the product library rejects the laboratory ones (SGLK_E_VARIANT).
static const Variant kVariants[] = {
    {"n128_s6_cta1_tma", 128, 6, 1, true, false, true},
    {"n128_s6_cta1_direct", 128, 6, 1, false, false, true},
    {"n256_s4_cta1_tma", 256, 4, 1, true, false, true},
    {"n256_s6_cta2_tma", 256, 6, 2, true, false, true},
    {"n256_s6_cta2_direct", 256, 6, 2, false, false, true},
    {"n128_s8_cta2_tma", 128, 8, 2, true, false, true},
    {"n64_s8_cta1_direct", 64, 8, 1, false, false, true},
    {"diag_n256_s6_cta2_nopromote", 256, 6, 2, false, false, false},   // diag_*: wrong results by design (pipeline ceilings)
    {"diag_n256_s6_cta2_plain_gemm", 256, 6, 2, true, false, false},
    {"diag_n256_s7_cta2_plain_gemm", 256, 7, 2, false, false, false},
    {"diag_n256_s4_cta2_plain_gemm", 256, 4, 2, false, false, false},
    {"diag_n256_s7_cta2_nopromote", 256, 7, 2, false, false, false},
    {"diag_n256_s6_cta2_plain_gemm_splitN", 256, 6, 2, true, false, false},
    {"n256_s6_cta2_tma_splitn", 256, 6, 2, true, false, true},
    {"n256_s4_cta1_tma_splitn", 256, 4, 1, true, false, true},
    {"diag_n256_s6_cta2_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_splitn_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_nopromote_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_ldtm_only_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_fma_only_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_fma_only_not_on_mma_scheduler_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_handoff_then_fma_timestamps", 256, 6, 2, true, false, false},
    {"diag_n256_s6_cta2_serialised_timestamps", 256, 6, 2, true, false, false},
    {"n256_s4_cl2_mcb_tma", 256, 4, 2, true, true, true},          // 2-CTA cluster, cta_group::1 tiles, multicast B
    {"lab_n256_s6_cta2_tma_pairld", 256, 6, 2, true, false, false},         // correct, measured no better: laboratory only
    {"lab_n256_s6_cta2_tma_splitn_pairld", 256, 6, 2, true, false, false},
    {"lab_n256_s4_cl2_mcb_tma_pairld", 256, 4, 2, true, true, false},
    {"lab_n128_s8_cta2_tma_earlyrel", 128, 8, 2, true, false, false},
    {"diag_n128_s8_cta2_nopromote", 128, 8, 2, true, false, false},
    {"n256_s5_cta2_tma_splitn_2box", 256, 5, 2, true, false, true},  // split-N pair tiles, two staging boxes per epilogue warp
    {"swapab_cta1", 128, 0, 1, false, false, true},                  // decode (gemm_decode.cu): weights are the MMA M side
    {"swapab_cta2", 128, 0, 2, false, false, true},                  // decode, CTA pair: 256 weight rows, tokens split over the pair
    {"swapab_rows", 128, 0, 1, false, false, true},                  // decode, row-balanced: one contiguous weight-row range per CTA, no split-K
    {"lab_rows_tokmajor", 128, 0, 1, false, false, false},           // laboratory: row-balanced with the tokens on the MMA M side (gemm_trows.cu);
                                                                     // correct, but epilogue-bound with two 256-column TMEM stages: 58 us against the
                                                                     // tile kernel's 41 at 128 x 24576 x 7168
};
constexpr int kDecodeVariant1 = 30, kDecodeVariant2 = 31, kRowsVariant = 32, kTRowsVariant = 33;
constexpr int kRowsMaxTokens = 64;     // the default dispatch uses the row-balanced swap-AB kernel up to here (pick_variant)
constexpr int kFirstStampVariant = 15, kLastStampVariant = 22;
constexpr int kNumVariants = int(sizeof(kVariants) / sizeof(kVariants[0]));

template <int BLOCK_N, int kStages, int kCG, typename OutT, bool kTmaStore, int kDiag = 0, bool kSplitN = false,
          bool kGrouped = false, int kEpi = 0, bool kMcB = false, int kPipe = 0, bool kWaitA = false>
static int launch_variant(const CUtensorMap& ta, const CUtensorMap& tb, const CMaps& tc, const GemmParams& p,
                          int workers, cudaStream_t st) {
  using C = Cfg<BLOCK_N, kCG>;
  auto kern = fp8_blockwise_gemm_kernel<BLOCK_N, kStages, kCG, OutT, kTmaStore, kDiag, kSplitN, kGrouped, kEpi, kMcB, kPipe, kWaitA>;
  constexpr int smem = kMaxDynSmem;
  (void)sizeof(C);
  static std::atomic<uint64_t> configured{0};   // per instantiation, one bit per device ordinal
  if (int e = ensure_max_dyn_smem(kern, configured, smem)) return e;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  constexpr int kCl = kMcB ? 2 : kCG;
  cfg.gridDim = dim3(unsigned(workers * kCl));
  cfg.blockDim = dim3(kNumThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return int(cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, p));
}

template <typename OutT>
static int launch_by_id(int v, const CUtensorMap& ta, const CUtensorMap& tb, const CMaps& tc,
                        const GemmParams& p, int num_sms, cudaStream_t st) {   // num_sms == workers here
  switch (v) {
    case 0: return launch_variant<128, 6, 1, OutT, true>(ta, tb, tc, p, num_sms, st);
    case 1: return launch_variant<128, 6, 1, OutT, false>(ta, tb, tc, p, num_sms, st);
    case 2: return launch_variant<256, 4, 1, OutT, true>(ta, tb, tc, p, num_sms, st);
    case 3: return launch_variant<256, 6, 2, OutT, true>(ta, tb, tc, p, num_sms, st);
    case 4: return launch_variant<256, 6, 2, OutT, false>(ta, tb, tc, p, num_sms, st);
    case 5: return launch_variant<128, 8, 2, OutT, true>(ta, tb, tc, p, num_sms, st);
    case 6: return launch_variant<64, 8, 1, OutT, false>(ta, tb, tc, p, num_sms, st);
    case 13: return launch_variant<256, 6, 2, OutT, true, 0, true>(ta, tb, tc, p, num_sms, st);
    case 14: return launch_variant<256, 4, 1, OutT, true, 0, true>(ta, tb, tc, p, num_sms, st);
    case 23: return launch_variant<256, 4, 1, OutT, true, 0, false, false, 0, true>(ta, tb, tc, p, num_sms, st);
    case 29: return launch_variant<256, 5, 2, OutT, true, 0, true, false, 0, false, 3>(ta, tb, tc, p, num_sms, st);
#ifdef SGLK_DIAG   // laboratory build only (libsglk_diag.so): pipeline-ceiling diagnostics with WRONG results, timeline stamps, retired experiments
    case 7: return launch_variant<256, 6, 2, OutT, false, 1>(ta, tb, tc, p, num_sms, st);
    case 8: return launch_variant<256, 6, 2, OutT, true, 2>(ta, tb, tc, p, num_sms, st);
    case 9: return launch_variant<256, 7, 2, OutT, false, 2>(ta, tb, tc, p, num_sms, st);
    case 10: return launch_variant<256, 4, 2, OutT, false, 2>(ta, tb, tc, p, num_sms, st);
    case 11: return launch_variant<256, 7, 2, OutT, false, 1>(ta, tb, tc, p, num_sms, st);
    case 12: return launch_variant<256, 6, 2, OutT, true, 3>(ta, tb, tc, p, num_sms, st);
    case 15: return launch_variant<256, 6, 2, OutT, true, 4>(ta, tb, tc, p, num_sms, st);
    case 16: return launch_variant<256, 6, 2, OutT, true, 4, true>(ta, tb, tc, p, num_sms, st);
    case 17: return launch_variant<256, 6, 2, OutT, true, 5>(ta, tb, tc, p, num_sms, st);
    case 18: return launch_variant<256, 6, 2, OutT, true, 6>(ta, tb, tc, p, num_sms, st);
    case 19: return launch_variant<256, 6, 2, OutT, true, 7>(ta, tb, tc, p, num_sms, st);
    case 20: return launch_variant<256, 6, 2, OutT, true, 8>(ta, tb, tc, p, num_sms, st);
    case 21: return launch_variant<256, 6, 2, OutT, true, 9>(ta, tb, tc, p, num_sms, st);
    case 22: return launch_variant<256, 6, 2, OutT, true, 10>(ta, tb, tc, p, num_sms, st);
    case 24: return launch_variant<256, 6, 2, OutT, true, 0, false, false, 0, false, 1>(ta, tb, tc, p, num_sms, st);
    case 25: return launch_variant<256, 6, 2, OutT, true, 0, true, false, 0, false, 1>(ta, tb, tc, p, num_sms, st);
    case 26: return launch_variant<256, 4, 1, OutT, true, 0, false, false, 0, true, 1>(ta, tb, tc, p, num_sms, st);
    case 27: return launch_variant<128, 8, 2, OutT, true, 0, false, false, 0, false, 2>(ta, tb, tc, p, num_sms, st);
    case 28: return launch_variant<128, 8, 2, OutT, true, 1>(ta, tb, tc, p, num_sms, st);
#endif
    default: return SGLK_E_VARIANT;
  }
}

static bool variant_available(int v) {
  if (v < 0 || v >= kNumVariants) return false;
#ifdef SGLK_DIAG
  return true;
#else
  return kVariants[v].product;
#endif
}

// fused output quantisation: the paired 256x256 tiles for prefill, the single-CTA 128x256 tiles below that
// The silu(gate)*up mode (kEpi == 2) is a LABORATORY form: correct (bit-identical to GEMM -> v2 quant op) but slower than
// the separate ops (3.26 against 2.07 ms at 16384 x 18432 x 7168 — the exact expf + IEEE division of silu_mul is a ~60-
// instruction dependent chain per element, run by 4 of the 12 warps while the tensor pipe waits).  It is compiled only
// into libsglk_diag.so; the product library answers SGLK_E_VARIANT.
template <typename OutT>
static int launch_quant_epilogue(int v, int mode, const CUtensorMap& ta, const CUtensorMap& tb, const CMaps& tc,
                                 const GemmParams& p, int workers, cudaStream_t st) {
#ifdef SGLK_DIAG
  if (mode == 2) {
    if (v == 13) return launch_variant<256, 6, 2, OutT, true, 0, true, false, 2>(ta, tb, tc, p, workers, st);
    if (v == 2) return launch_variant<256, 4, 1, OutT, true, 0, false, false, 2>(ta, tb, tc, p, workers, st);
    return SGLK_E_VARIANT;
  }
#endif
  if (mode != 1) return SGLK_E_VARIANT;
  if (v == 13) return launch_variant<256, 6, 2, OutT, true, 0, true, false, 1>(ta, tb, tc, p, workers, st);   // split-N pair tiles
  if (v == 2) return launch_variant<256, 4, 1, OutT, true, 0, false, false, 1>(ta, tb, tc, p, workers, st);
  return SGLK_E_VARIANT;
}

static bool tma_store_ok(const void* C, int64_t ldc, int out_bytes) {
  return (reinterpret_cast<uintptr_t>(C) & 15) == 0 && (ldc * out_bytes) % 16 == 0;
}

// Prefill default: 29 (split-N pair tiles, two staging boxes per epilogue warp, 5-stage ring) — +1-3 % over 13 at every
// prefill shape, whole GPU suite green on it.  Decode (M <= 256, plain [M, N] output): the swap-AB kernel of
// gemm_decode.cu, single CTAs for a few tokens, CTA pairs (tokens split over the pair: half the L2 -> smem traffic of the
// activation tile) from 33 tokens up.
static int pick_variant(int64_t M, int64_t N, int64_t K, bool plain_output, int num_sms = 0) {
  // Row-balanced swap-AB kernel (gemm_decode.cu, swapab_rows): every CTA streams one contiguous range of weight rows —
  // R = ceil8(N / #SMs) rows for all of K, or for narrow weights R = ceil8(N / #clusters) rows for a half / a quarter of K
  // with the cluster's partials summed through distributed shared memory.  Measured in a CUDA graph on B200 with every
  // launch streaming its weight from HBM (tools/decode_ab.py, K=7168, us; best tile kernel = min of the swap-AB tiles
  // 30/31 and the 128-row token-major tiles 0):
  //      M =          1      16      32      64      128
  //   N=24576 rows  29.5    29.7    30.9    32.8    49.6        (0.91 .. 0.84 of the measured copy bandwidth up to 64)
  //           tiles 36.1    36.6    37.1    38.1    40.8
  //   N=2112  rows   8.6     9.1    10.0    12.7    17.6        (clusters of 4 over K)
  //           tiles 11.7    11.7    13.2    15.0    15.4
  //   N=7168  rows    -     13.9      -     16.5    20.7        (clusters of 2)
  //           tiles   -     16.1      -     19.0    20.6
  // Beyond 64 tokens every CTA re-reading its K range of the activations from L2 costs more than the balance gains.
  if (plain_output && M <= kRowsMaxTokens && rows_supported(M, N, K)) return kRowsVariant;
  // 65..256 tokens are still weight streams; what matters is how the tile count fits the SMs (tools/decode_ab.py, K=7168,
  // GPU-side in a CUDA graph, us):
  //   M=128: N=24576: 128x256 single-CTA split-N tiles (14) 36.1 — 96 tiles, ONE wave — against 40.9 for 128x128 tiles
  //          (0: 192 tiles, stream-K) and 48.4 for pair tiles (29); N=7168: 0 19.4-20.7, 14 22.5-24.6 (28 tiles only)
  //   M=256: N=7168: 256x128 pair tiles (5) 19.8 — 56 tiles, one wave — against 27.5 (29) and 21.2 (0);
  //          N=24576: 29 52.5, 5 55.7, 0 71.2; N=2112: 0 16.4, 5 16.5, 29 22.2
  const int64_t sms = num_sms > 0 ? num_sms : device_sm_count();   // the current device's count (148 on B200; also the answer without a device)
  if (plain_output && M > kRowsMaxTokens && M <= 128) {
    const int64_t t256 = (N + 255) / 256;
    if (t256 <= sms && 5 * t256 >= 3 * sms) return 14;      // one wave of 256-wide tiles that fills >= 60 % of the SMs
  }
  // a handful of wide tiles on a narrow weight (N=2112, 129 <= M <= 256: 9 pair tiles for 74 pairs): 128-wide single-CTA
  // tiles give stream-K twice the units to balance
  if (M > 128 && M <= 256 && N < 4096) return 0;
  if (plain_output && M > 128 && M <= 256 && (N + 127) / 128 <= sms / 2) return 5;   // one wave of 256x128 pair tiles
  // paired 256x256 tiles (one pair covers 256 rows), each K block issued as two 128-column MMA halves with their own
  // hand-off barriers: a half's stage has three half-MMAs (768 clk) of slack before it is needed again instead of one
  // whole MMA (512) — +5-10 % over the unsplit pair tile at every prefill shape measured
  if (M > 128 && N >= 256) return 29;
  // M <= 128 without the decode kernel: one m-block; 128-wide tiles, balanced over the SMs by stream-K
  return 0;
}

struct Plan {
  int variant, workers, stream_k;
  int64_t scale_bytes;     // workspace for re-laid-out scales (0: not worth it / not possible)
  int32_t num_m_blocks, num_n_blocks, num_k_blocks;
  int64_t counter_bytes, partial_bytes;
};

// Tile shape, worker count and whether to cut the work stream-K style (needs a workspace).  Cost model in units of
// one K block of one tile: whole tiles cost waves * G per worker; stream-K costs ceil(tiles * G / workers) plus the
// fixup — partials written, fenced and re-read by the last arriver while the MMA pipe idles (the accumulators live
// in the epilogue warps' registers).  Measured on B200 the fixup is ~15 us for the 256-wide pair tiles (128 KB of
// fp32 partial per CTA), i.e. ~30 K blocks; it scales with the partial size.  Decode shapes win clearly (N=24576:
// 1.3 waves of 128-wide tiles, N=2112: 17 tiles for 148 SMs); short-K prefill shapes (4096x7168x2048: 75.8 us
// stream-K vs 67.6 us whole tiles) do not and stay on whole tiles.
// flags (SGLK_GEMM_* of the C ABI): stream-K decision (cost model / whenever possible / never), raster group override
static Plan make_plan(int64_t M, int64_t N, int64_t K, int variant, int num_sms, bool have_ws, int flags) {
  const int sk_mode = (flags & SGLK_GEMM_STREAMK_FORCE) ? 1 : ((flags & SGLK_GEMM_STREAMK_NEVER) ? -1 : 0);
  Plan pl;
  pl.variant = variant;
  const Variant v = kVariants[variant];
  pl.num_m_blocks = int32_t((M + BLOCK_M * v.cg - 1) / (BLOCK_M * v.cg));
  pl.num_n_blocks = int32_t((N + v.block_n - 1) / v.block_n);
  pl.num_k_blocks = int32_t((K + BLOCK_K - 1) / BLOCK_K);
  const int64_t tiles = int64_t(pl.num_m_blocks) * pl.num_n_blocks;
  int w0 = num_sms / v.cg;
  if (w0 < 1) w0 = 1;
  const int64_t waves = (tiles + w0 - 1) / w0;
  const double eff = double(tiles) / double(waves * w0);
  const int64_t units = tiles * pl.num_k_blocks;
  const int64_t sk_workers = tiles * 4 < w0 ? tiles * 4 : w0;
  const int64_t sk_cost = (units + sk_workers - 1) / sk_workers + 30 * v.block_n / 256;
  const bool possible = have_ws && (kVariants[variant].product || variant >= 13) && pl.num_k_blocks >= 2 && eff < 0.9 && units >= w0;
  pl.stream_k = (possible && sk_mode >= 0 && (sk_mode > 0 || sk_cost < waves * pl.num_k_blocks)) ? 1 : 0;
  pl.workers = int(pl.stream_k ? w0 : (tiles < w0 ? tiles : w0));
  // few, large tiles (N=2112: 17): more than ~4 contributors per tile makes the last arriver's sequential
  // re-read of the partials longer than the K blocks it saves (measured: 68 workers 13.6 us, 148 workers 18.9 us)
  if (pl.stream_k && tiles * 4 < pl.workers) pl.workers = int(tiles * 4);
  if (pl.workers < 1) pl.workers = 1;
  // the counter area has a FIXED size so that launches with different tilings can share one workspace:
  // partials of one launch must never land where another launch expects zeroed counters
  pl.counter_bytes = kCounterAreaBytes;
  if (tiles * v.cg * kNumEpiWarps * 4 > kCounterAreaBytes) { pl.stream_k = 0; pl.workers = int(tiles < w0 ? tiles : w0); }
  pl.partial_bytes = int64_t(2) * pl.workers * v.cg * kNumEpiWarps * 32 * (v.block_n / 2) * 4;
  // prefill-sized problems: scales that are not K-block-contiguous are re-laid-out once per call (workspace permitting)
  const int64_t nb128 = (N + 127) / 128;
  // (measured: +9 % at 16384x18432x7168, a loss below ~100 us of GEMM because of the extra launch)
  pl.scale_bytes = (have_ws && M >= 512 && double(M) * double(N) * double(K) >= 1.5e11 && pl.num_k_blocks % 4 == 0 && v.block_n >= 128)
                       ? ((M * pl.num_k_blocks * 4 + 255) / 256 * 256 + (nb128 * pl.num_k_blocks * 4 + 255) / 256 * 256) : 0;
  return pl;
}

// Fused output quantisation request (sglk_fp8_blockwise_scaled_mm_quant): codes [M, Nq] (Nq = N, or N/2 with
// silu_mul), row stride ldq bytes; scales (row, group) at row * ss_m + group * ss_g floats.
struct QuantOut {
  void* codes;
  int64_t ldq;
  float* scales;
  int64_t ss_m, ss_g;
  float eps;
  int silu_mul;
};

// The gated kernel's walk over the m-blocks (GemmParams::a_ready): num_m_blocks, num_n_blocks and group_m are set.
static void set_panel_walk(GemmParams& p, int rank, int world, int whole_panels) {
  p.panel_world = world;
  p.panel_rank = rank;
  p.panel_own = p.num_m_blocks / world;
  if (p.group_m > p.panel_own) p.group_m = p.panel_own;
  const int w1 = world - 1;
  p.panel_whole = whole_panels ? 1 : 0;
  // rasterisation groups of the remote part: whole slices (all-at-once arrival) or up to 8 m-blocks of one panel
  p.panel_group = p.panel_whole ? (p.panel_own < 8 ? p.panel_own : 8) : (w1 > 0 ? w1 * (8 / w1 > 0 ? 8 / w1 : 1) : 1);
}

// Gated request (sglk_fp8_blockwise_scaled_mm_gated): one arrival counter per 256 rows of A, see GemmParams::a_ready.
struct GateIn {
  const uint32_t* counters;
  uint32_t target;
  int rank, world;
  int whole_panels;   // arrival order of the other ranks' panels: 0 = all at once, slice by slice; 1 = one panel after the other
  uint32_t wait_polls; // see GemmParams::a_wait_polls
};

static int gemm_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M,
                      int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk,
                      int64_t sb_sk, int64_t sb_sn, int out_dtype, int variant, int num_sms, int flags, void* workspace,
                      int64_t workspace_bytes, void* stream, void* const* extra_dests = nullptr, int num_extra = 0,
                      void* mc_ptr = nullptr, const QuantOut* qo = nullptr, const GateIn* gate = nullptr) {
  if (M < 0 || N < 0 || K < 0) return SGLK_E_SHAPE;
  if (out_dtype != SGLK_BF16 && out_dtype != SGLK_F16) return SGLK_E_DTYPE;
  if (M == 0 || N == 0) return SGLK_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (qo) {
    // whole 1x128 groups only; gate|up tiles are 256 wide; codes leave through a 128-byte-wide TMA box
    if (!qo->codes || !qo->scales) return SGLK_E_NULL;
    if (K == 0 || N % (qo->silu_mul ? 256 : 128) != 0) return SGLK_E_SHAPE;
    const int64_t nq = qo->silu_mul ? N / 2 : N;
    if (qo->ldq < nq || qo->ldq % 16 != 0 || (reinterpret_cast<uintptr_t>(qo->codes) & 15)) return SGLK_E_STRIDE;
    if (num_extra != 0 || mc_ptr) return SGLK_E_ARG;
    variant = M > 128 ? 13 : 2;
    Cout = qo->codes; ldc = N;       // (placeholders for the shared checks below; C itself is never written)
  }
  if (!Cout) return SGLK_E_NULL;
  if (K == 0) {   // empty sum: C = 0
    if (ldc == N) return int(cudaMemsetAsync(Cout, 0, size_t(M) * N * 2, st));
    return int(cudaMemset2DAsync(Cout, size_t(ldc) * 2, 0, size_t(N) * 2, size_t(M), st));
  }
  if (!A || !B || !sa || !sb) return SGLK_E_NULL;
  if (K % 16 != 0 || lda % 16 != 0 || ldb % 16 != 0 || lda < K || ldb < K || ldc < N) return SGLK_E_STRIDE;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return SGLK_E_STRIDE;
  if (M >= (int64_t(1) << 31) || N >= (int64_t(1) << 31) || K >= (int64_t(1) << 31)) return SGLK_E_SHAPE;
  const bool plain_output = !qo && num_extra == 0 && !mc_ptr && !gate;
  if (gate) {
    // the paired 256-row tiles, one counter per m-block, equal whole panels per rank, whole tiles (no stream-K)
    if (!gate->counters || gate->world < 1 || gate->rank < 0 || gate->rank >= gate->world) return SGLK_E_ARG;
    if (M % (int64_t(2 * BLOCK_M) * gate->world) != 0 || N < 256 || qo || num_extra != 0 || mc_ptr) return SGLK_E_SHAPE;
    variant = 29;
    flags = (flags & ~3) | SGLK_GEMM_STREAMK_NEVER;
  }
  const bool picked = variant < 0;
  if (picked) variant = pick_variant(M, N, K, plain_output, num_sms);
  if (!variant_available(variant)) return SGLK_E_VARIANT;
#ifdef SGLK_DIAG
  if (variant == kTRowsVariant) {
    if (!plain_output || !trows_supported(M, N, K)) return SGLK_E_VARIANT;
    return trows_entry(A, B, sa, sb, Cout, M, N, K, lda, ldb, ldc, sa_sm, sa_sk, sb_sk, sb_sn, out_dtype, num_sms, st);
  }
#endif
  if (variant == kRowsVariant) {
    if (!plain_output || !rows_supported(M, N, K)) return SGLK_E_VARIANT;
    const int rc = rows_entry(A, B, sa, sb, Cout, M, N, K, lda, ldb, ldc, sa_sm, sa_sk, sb_sk, sb_sn, out_dtype, num_sms, flags, st);
    // A kernel the HEURISTIC chose must not surface its own planning limits to the user: if the row-balanced kernel
    // cannot be planned for this problem (no ring that fits, no launch configuration, tensor map refused — all found
    // before anything is launched), the general 128x128 tile kernel below takes it.  An explicitly requested variant
    // still reports the error.
    if (!picked || (rc != SGLK_E_SHAPE && rc != SGLK_E_VARIANT && rc != SGLK_E_DRIVER)) return rc;
    variant = 0;
  }
  if (variant == kDecodeVariant1 || variant == kDecodeVariant2) {
    if (!plain_output || !decode_supported(M, N, K)) return SGLK_E_VARIANT;
    return decode_entry(A, B, sa, sb, Cout, M, N, K, lda, ldb, ldc, sa_sm, sa_sk, sb_sk, sb_sn, out_dtype,
                        variant == kDecodeVariant2 ? 2 : 1, num_sms, flags, workspace, workspace_bytes, st);
  }
  Variant v = kVariants[variant];
  if (num_extra < 0 || num_extra > kMaxDests - 1 || (num_extra > 0 && !extra_dests)) return SGLK_E_ARG;
  for (int d = 0; d < num_extra; ++d)
    if (!extra_dests[d] || !tma_store_ok(extra_dests[d], ldc, 2)) return SGLK_E_STRIDE;
  if (!qo && v.tma_store && !tma_store_ok(Cout, ldc, 2)) {
    if (gate) return SGLK_E_STRIDE;
    // same tile shape, predicated-store epilogue
    variant = (v.cg == 2) ? 4 : 1;
    v = kVariants[variant];
  }
  if (num_sms <= 0) num_sms = device_sm_count();

  GemmParams p;
  p.grp_tiles = nullptr; p.grp_num_m_tiles = nullptr; p.grp_sb_expert_stride = 0;
  p.a_ready = nullptr; p.a_ready_target = 0; p.panel_own = p.panel_rank = 0; p.panel_world = 1; p.panel_group = 1; p.panel_whole = 0;
  p.sa = sa; p.sb = sb; p.C = Cout;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.sa_sm = sa_sm; p.sa_sk = sa_sk; p.sb_sk = sb_sk; p.sb_sn = sb_sn;
  // gate|up mode pairs two epilogue warps with named barriers: both must finish the same tiles, so no stream-K there
  const bool allow_sk = workspace != nullptr && !(qo && qo->silu_mul);
  Plan pl = make_plan(M, N, K, variant, num_sms, allow_sk, flags);
  if (qo) { p.q_scales = qo->scales; p.qs_sm = qo->ss_m; p.qs_sg = qo->ss_g; p.q_eps = qo->eps; }
  else { p.q_scales = nullptr; p.qs_sm = p.qs_sg = 0; p.q_eps = 0.0f; }
  if (pl.stream_k && (workspace_bytes < pl.counter_bytes + pl.partial_bytes || (reinterpret_cast<uintptr_t>(workspace) & 15)))
    pl = make_plan(M, N, K, variant, num_sms, false, flags);
  p.num_m_blocks = pl.num_m_blocks;
  p.num_n_blocks = pl.num_n_blocks;
  p.num_k_blocks = pl.num_k_blocks;
  p.num_sb_cols = int32_t((N + 127) / 128);
  // Rasterisation group: its A panels (group_m * tile_m * K bytes) must stay in L2 while B streams past once per group.
  // ncu at 16384x18432x7168 (tools/session22.sh): DRAM reads 5.60 / 3.74 / 1.65 / 1.03 / 1.70 / 4.42 GB for groups of
  // 2 / 4 / 8 / 16 / 32 / 64 m-blocks (algorithmic 0.25 GB) at equal speed within noise — ~32 MB of A is what one L2
  // partition keeps; beyond that the panels evict each other.
  {
    int64_t gm = (int64_t(32) << 20) / (int64_t(BLOCK_M) * v.cg * K);
    gm = gm < 4 ? 4 : (gm > 16 ? 16 : gm);
    const int gm_override = (flags >> SGLK_GEMM_GROUP_M_SHIFT) & 0xFF;
    p.group_m = gm_override > 0 ? gm_override : int32_t(gm);
  }
  p.stream_k = pl.stream_k;
  p.ws_counters = pl.stream_k ? static_cast<int32_t*>(workspace) : nullptr;
  p.ws_partials = pl.stream_k ? reinterpret_cast<float*>(static_cast<char*>(workspace) + pl.counter_bytes) : nullptr;
  {
    const bool vec_layout = sa_sk == 1 && sb_sk == 1 && sa_sm % 4 == 0 && sb_sn % 4 == 0 &&
                            (reinterpret_cast<uintptr_t>(sa) & 15) == 0 && (reinterpret_cast<uintptr_t>(sb) & 15) == 0;
    const int64_t base = pl.stream_k ? pl.counter_bytes + pl.partial_bytes : kCounterAreaBytes;
    if (gate && !vec_layout) return SGLK_E_STRIDE;   // the scales arrive during the launch: they cannot be re-laid-out first
    if (!vec_layout && pl.scale_bytes > 0 && workspace && workspace_bytes >= base + pl.scale_bytes &&
        (reinterpret_cast<uintptr_t>(workspace) & 15) == 0 && (variant < kFirstStampVariant || variant > kLastStampVariant)) {
      const int G = pl.num_k_blocks, NB = p.num_sb_cols;
      float* sa_t = reinterpret_cast<float*>(static_cast<char*>(workspace) + base);
      float* sb_t = sa_t + (int64_t(M) * G * 4 + 255) / 256 * 64;
      // one launch: activation scales (m, kb) -> [m][kb], weight scales (kb, nb) -> [nb][kb]
      const RelayoutJob ja{sa, sa_t, sa_sm, sa_sk, int(M), G, int((M + 31) / 32)};
      const RelayoutJob jb{sb, sb_t, sb_sn, sb_sk, NB, G, (NB + 31) / 32};
      const int tiles_a = ja.tiles_r * ((G + 31) / 32), tiles_b = jb.tiles_r * ((G + 31) / 32);
      relayout_scales_kernel<<<unsigned(tiles_a + tiles_b), 256, 0, st>>>(ja, jb, tiles_a);
      p.sa = sa_t; p.sa_sm = G; p.sa_sk = 1;
      p.sb = sb_t; p.sb_sn = G; p.sb_sk = 1;
      sa = sa_t; sb = sb_t; sa_sm = G; sa_sk = 1; sb_sn = G; sb_sk = 1;
    }
  }
  p.dbg = nullptr;
  if (variant >= kFirstStampVariant && variant <= kLastStampVariant) {   // timestamps diagnostic: the stamps go to the start of the partials area of the workspace
    if (!workspace || workspace_bytes < kCounterAreaBytes + 64 * 8 * 8) return SGLK_E_ARG;
    p.dbg = reinterpret_cast<long long*>(static_cast<char*>(workspace) + kCounterAreaBytes);
  }

  // Decode (M < 128): only load the rows that exist.  TMA zero-fills out-of-bounds rows of a box at
  // ~20 clk per row (measured: M=1 ran 2x slower than M=128), so the A box is cut to round_up(M, 8) rows;
  // the MMA still reads 128 rows of smem, the stale ones only feed accumulator lanes that are never stored.
  const uint32_t a_rows = (v.cg == 1 && M < BLOCK_M) ? uint32_t((M + 7) / 8 * 8) : uint32_t(BLOCK_M);
  // bytes of B in one CTA's stage / bytes its full barrier sees per stage (pair: both CTAs' loads on the leader's barrier;
  // multicast-B: own A + own half of B + the peer's half)
  const uint32_t b_bytes = uint32_t(v.mcb ? v.block_n : v.block_n / v.cg) * BLOCK_K;
  p.stage_tx_bytes = v.mcb ? (a_rows * BLOCK_K + b_bytes) : uint32_t(v.cg) * (a_rows * BLOCK_K + b_bytes);
  p.a_slot_bytes = (a_rows * BLOCK_K + 1023u) & ~1023u;
  p.stage_bytes = p.a_slot_bytes + b_bytes;
  {
    // what follows the ring: the TMA-store staging boxes, or (direct store) nothing — but the MMA reads a
    // full 16 KB of A rows from every stage, so the last stage needs that much addressable smem behind it
    const int tail = v.tma_store ? (variant == 29 ? 65536 : 32768) : 16384;
    int n = (kMaxDynSmem - kSmemOverhead - tail) / int(p.stage_bytes);
    if (n > kMaxStages) n = kMaxStages;
    if (n > v.stages && a_rows == uint32_t(BLOCK_M)) n = v.stages;   // prefill: the tuned depth of the variant
    if (n < 2) return SGLK_E_SHAPE;
    p.num_stages = n;
  }
  CUtensorMap ta, tb;
  CMaps tc;
  p.num_dests = 1 + num_extra;
  p.mc_ptr = mc_ptr;
  p.scale_vec = (sa_sk == 1 && sb_sk == 1 && p.num_k_blocks % 4 == 0 && sa_sm % 4 == 0 && sb_sn % 4 == 0 &&
                 (reinterpret_cast<uintptr_t>(sa) & 15) == 0 && (reinterpret_cast<uintptr_t>(sb) & 15) == 0) ? 1 : 0;
  if (mc_ptr && (!v.tma_store || (reinterpret_cast<uintptr_t>(mc_ptr) & 15) || N % 8 != 0 || ldc % 8 != 0)) return SGLK_E_STRIDE;
  if (!make_tmap_2d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, A, uint64_t(M), uint64_t(K), uint64_t(lda), a_rows, BLOCK_K))
    return SGLK_E_DRIVER;
  const bool split_b_boxes = (variant == 13 || variant == 16 || variant == 25 || variant == 29);   // pair + split-N: two 64-row boxes per CTA per stage
  if (!make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(N), uint64_t(K), uint64_t(ldb),
                    uint32_t(v.block_n / v.cg / (split_b_boxes ? 2 : 1)), BLOCK_K))
    return SGLK_E_DRIVER;
  if (qo) {
    const int64_t nq = qo->silu_mul ? N / 2 : N;
    if (!make_tmap_2d(&tc.m[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, qo->codes, uint64_t(M), uint64_t(nq), uint64_t(qo->ldq), 32, 128))
      return SGLK_E_DRIVER;
    for (int d = 1; d < kMaxDests; ++d) tc.m[d] = tc.m[0];
    const int mode = qo->silu_mul ? 2 : 1;
    if (out_dtype == SGLK_BF16) return launch_quant_epilogue<__nv_bfloat16>(variant, mode, ta, tb, tc, p, pl.workers, st);
    return launch_quant_epilogue<__half>(variant, mode, ta, tb, tc, p, pl.workers, st);
  }
  if (v.tma_store) {
    const CUtensorMapDataType cdt = out_dtype == SGLK_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    if (!make_tmap_2d(&tc.m[0], cdt, 2, Cout, uint64_t(M), uint64_t(N), uint64_t(ldc), 32, kStoreBoxCols)) return SGLK_E_DRIVER;
    for (int d = 0; d < num_extra; ++d)
      if (!make_tmap_2d(&tc.m[1 + d], cdt, 2, extra_dests[d], uint64_t(M), uint64_t(N), uint64_t(ldc), 32, kStoreBoxCols))
        return SGLK_E_DRIVER;
    for (int d = p.num_dests; d < kMaxDests; ++d) tc.m[d] = tc.m[0];
  } else {
    if (num_extra > 0) return SGLK_E_STRIDE;   // multi-destination output needs the TMA-store epilogue
    for (int d = 0; d < kMaxDests; ++d) tc.m[d] = ta;   // unused by the kernel
  }
  if (gate) {
    if (!p.scale_vec) return SGLK_E_STRIDE;
    p.a_ready = gate->counters;
    p.a_ready_target = gate->target;
    p.a_wait_polls = gate->wait_polls;
    set_panel_walk(p, gate->rank, gate->world, gate->whole_panels);
    if (out_dtype == SGLK_BF16)
      return launch_variant<256, 5, 2, __nv_bfloat16, true, 0, true, false, 0, false, 3, true>(ta, tb, tc, p, pl.workers, st);
    return launch_variant<256, 5, 2, __half, true, 0, true, false, 0, false, 3, true>(ta, tb, tc, p, pl.workers, st);
  }
  if (out_dtype == SGLK_BF16) return launch_by_id<__nv_bfloat16>(variant, ta, tb, tc, p, pl.workers, st);
  return launch_by_id<__half>(variant, ta, tb, tc, p, pl.workers, st);
}

// Grouped (MoE) form: experts' activations stacked along M, one [N, K] weight per expert, m-tiles from a device table.
static int grouped_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout,
                         const int32_t* expert_offsets, const int32_t* problem_sizes, int64_t num_experts,
                         int64_t total_m, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm,
                         int64_t sa_sk, int64_t sb_se, int64_t sb_sk, int64_t sb_sn, int out_dtype, int num_sms,
                         void* workspace, int64_t workspace_bytes, void* stream) {
  if (total_m < 0 || N < 0 || K < 0 || num_experts < 0) return SGLK_E_SHAPE;
  if (out_dtype != SGLK_BF16 && out_dtype != SGLK_F16) return SGLK_E_DTYPE;
  if (total_m == 0 || N == 0 || num_experts == 0) return SGLK_OK;
  if (!A || !B || !sa || !sb || !Cout || !expert_offsets || !problem_sizes || !workspace) return SGLK_E_NULL;
  if (K <= 0 || K % 16 != 0 || lda % 16 != 0 || ldb % 16 != 0 || lda < K || ldb < K || ldc < N) return SGLK_E_STRIDE;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
      (reinterpret_cast<uintptr_t>(workspace) & 15)) return SGLK_E_STRIDE;
  if (total_m >= (int64_t(1) << 31) || num_experts * N >= (int64_t(1) << 31) || K >= (int64_t(1) << 31)) return SGLK_E_SHAPE;
  const int64_t max_tiles = (total_m + BLOCK_M - 1) / BLOCK_M + num_experts;   // every expert can add one ragged tile
  if (workspace_bytes < 16 + max_tiles * 16) return SGLK_E_ARG;
  if (num_sms <= 0) num_sms = device_sm_count();
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int32_t* count = static_cast<int32_t*>(workspace);
  int4* table = reinterpret_cast<int4*>(static_cast<char*>(workspace) + 16);
  // Tile shape.  Prefill-sized expert batches (>= 256 tokens per expert on average, whole 256-column blocks, 16-byte
  // aligned output rows) take the dense prefill kernel's tiles: 256-row CTA-pair tiles, split-N hand-offs, two staging
  // boxes per epilogue warp — the finished boxes are read back row-contiguously and stored with the rows cut at the
  // expert's end.  Everything else: single-CTA 128-row tiles with per-thread predicated stores.
  const bool pair = (total_m / num_experts >= 256) && N % 256 == 0 && tma_store_ok(Cout, ldc, 2) && num_sms >= 2;
  const int tile_rows = pair ? 2 * BLOCK_M : BLOCK_M;
  build_group_tiles_kernel<<<1, 1024, 0, st>>>(expert_offsets, problem_sizes, int(num_experts), table, count, int(max_tiles),
                                               tile_rows);

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.sa = sa; p.sb = sb; p.C = Cout;
  p.M = total_m; p.N = N; p.K = K; p.ldc = ldc;
  p.sa_sm = sa_sm; p.sa_sk = sa_sk; p.sb_sk = sb_sk; p.sb_sn = sb_sn;
  p.num_m_blocks = int32_t(max_tiles);
  p.num_n_blocks = int32_t((N + 127) / 128);
  p.num_k_blocks = int32_t((K + BLOCK_K - 1) / BLOCK_K);
  p.num_sb_cols = int32_t((N + 127) / 128);
  p.group_m = 0;     // grouped mode: tile order switch (see grouped_tile in the kernel): n-block fastest
  p.num_dests = 1;
  p.grp_tiles = table; p.grp_num_m_tiles = count; p.grp_sb_expert_stride = sb_se;
  const uint32_t b_bytes = 128u * BLOCK_K;
  p.stage_tx_bytes = BLOCK_M * BLOCK_K + b_bytes;
  p.a_slot_bytes = BLOCK_M * BLOCK_K;
  p.stage_bytes = p.a_slot_bytes + b_bytes;
  p.num_stages = 6;
  p.scale_vec = (sa_sk == 1 && sb_sk == 1 && p.num_k_blocks % 4 == 0 && sa_sm % 4 == 0 && sb_sn % 4 == 0 && sb_se % 4 == 0 &&
                 (reinterpret_cast<uintptr_t>(sa) & 15) == 0 && (reinterpret_cast<uintptr_t>(sb) & 15) == 0) ? 1 : 0;
  CUtensorMap ta, tb;
  CMaps tc;
  if (!make_tmap_2d(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, A, uint64_t(total_m), uint64_t(K), uint64_t(lda), BLOCK_M, BLOCK_K))
    return SGLK_E_DRIVER;
  if (!make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(num_experts * N), uint64_t(K), uint64_t(ldb), 128, BLOCK_K))
    return SGLK_E_DRIVER;
  for (int d = 0; d < kMaxDests; ++d) tc.m[d] = ta;   // unused by the predicated-store epilogue
  int64_t workers = max_tiles * p.num_n_blocks;
  if (workers > num_sms) workers = num_sms;
  if (pair) {
    // operand ring as gemm_entry sets it up for variant 29: per CTA 128 rows of A and two 64-row boxes of B per stage
    p.num_n_blocks = int32_t(N / 256);
    const uint32_t bb = 128u * BLOCK_K;
    p.stage_tx_bytes = 2u * (BLOCK_M * BLOCK_K + bb);
    p.a_slot_bytes = BLOCK_M * BLOCK_K;
    p.stage_bytes = p.a_slot_bytes + bb;
    p.num_stages = 5;
    if (!make_tmap_2d(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, B, uint64_t(num_experts * N), uint64_t(K), uint64_t(ldb), 64, BLOCK_K))
      return SGLK_E_DRIVER;
    int64_t w = max_tiles * p.num_n_blocks;
    if (w > num_sms / 2) w = num_sms / 2;
    if (out_dtype == SGLK_BF16)
      return launch_variant<256, 5, 2, __nv_bfloat16, true, 0, true, true, 0, false, 3>(ta, tb, tc, p, int(w), st);
    return launch_variant<256, 5, 2, __half, true, 0, true, true, 0, false, 3>(ta, tb, tc, p, int(w), st);
  }
  if (out_dtype == SGLK_BF16)
    return launch_variant<128, 6, 1, __nv_bfloat16, false, 0, false, true>(ta, tb, tc, p, int(workers), st);
  return launch_variant<128, 6, 1, __half, false, 0, false, true>(ta, tb, tc, p, int(workers), st);
}

}  // namespace sglk

extern "C" int sglk_fp8_blockwise_scaled_grouped_mm(const void* A, const void* B, const float* sa, const float* sb, void* C,
                                                    const int32_t* expert_offsets, const int32_t* problem_sizes,
                                                    int64_t num_experts, int64_t total_m, int64_t N, int64_t K,
                                                    int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_stride_m,
                                                    int64_t sa_stride_k, int64_t sb_stride_e, int64_t sb_stride_k,
                                                    int64_t sb_stride_n, int out_dtype, int num_sms, void* workspace,
                                                    int64_t workspace_bytes, void* stream) {
  return sglk::grouped_entry(A, B, sa, sb, C, expert_offsets, problem_sizes, num_experts, total_m, N, K, lda, ldb, ldc,
                             sa_stride_m, sa_stride_k, sb_stride_e, sb_stride_k, sb_stride_n, out_dtype, num_sms,
                             workspace, workspace_bytes, stream);
}

extern "C" int64_t sglk_grouped_mm_workspace_bytes(int64_t num_experts, int64_t total_m) {
  if (num_experts < 0 || total_m < 0) return 0;
  return 16 + ((total_m + sglk::BLOCK_M - 1) / sglk::BLOCK_M + num_experts) * 16;
}

// --------------------------------------------------------------------------------------------
// Grouped GEMM for hosts that cannot allocate device memory (see include/sglk_fp8.h).
// The m-tile table of the grouped kernel normally lives in a caller-supplied workspace.  Here it lives in a fixed array
// that is part of the library image: every GPU that loads libsglk holds SGLK_GROUPED_STATIC_SCRATCH_BYTES of it until
// the process exits.  In HOST code the identifier below is only a handle for the runtime's symbol table — it is not a
// device address and must never be handed to a kernel; the address the GPU sees is asked for once per device with
// cudaGetSymbolAddress (each device has its own copy of the array at its own address).
// --------------------------------------------------------------------------------------------
namespace sglk {

static __device__ __align__(256) unsigned char g_grouped_static_scratch[SGLK_GROUPED_STATIC_SCRATCH_BYTES];

constexpr int kMaxStaticScratchDevices = 64;

// Device address of the CURRENT device's copy of the array.  The first call on a device asks the runtime (which may load
// the module there — do it outside stream capture); later calls read the cached pointer.  No allocation anywhere.
static int grouped_static_scratch(void** out) {
  static std::mutex mu;
  static void* addr[kMaxStaticScratchDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  if (dev < 0 || dev >= kMaxStaticScratchDevices) return SGLK_E_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_grouped_static_scratch);
    if (e != cudaSuccess) return int(e);
    if (!p) return int(cudaErrorInvalidSymbol);
    addr[dev] = p;
  }
  *out = addr[dev];
  return SGLK_OK;
}

}  // namespace sglk

extern "C" int64_t sglk_grouped_mm_static_max_tiles(void) {
  return (int64_t(SGLK_GROUPED_STATIC_SCRATCH_BYTES) - 16) / 16;
}

extern "C" int sglk_fp8_blockwise_scaled_grouped_mm_static(const void* A, const void* B, const float* sa, const float* sb,
                                                           void* C, const int32_t* expert_offsets,
                                                           const int32_t* problem_sizes, int64_t num_experts,
                                                           int64_t total_m, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                                                           int64_t ldc, int64_t sa_stride_m, int64_t sa_stride_k,
                                                           int64_t sb_stride_e, int64_t sb_stride_k, int64_t sb_stride_n,
                                                           int out_dtype, int num_sms, void* stream) {
  // capacity first, so an oversized problem is refused whether or not a GPU is present
  if (num_experts > 0 && total_m > 0 &&
      sglk_grouped_mm_workspace_bytes(num_experts, total_m) > int64_t(SGLK_GROUPED_STATIC_SCRATCH_BYTES))
    return SGLK_E_CAPACITY;
  void* scratch = nullptr;
  if (num_experts > 0 && total_m > 0 && N > 0) {   // empty problems launch nothing and need no scratch
    const int rc = sglk::grouped_static_scratch(&scratch);
    if (rc != SGLK_OK) return rc;
  }
  // Nothing in the array has to start at zero: the table-building kernel rewrites the tile count and every entry the
  // GEMM will read, on `stream`, before the GEMM runs.  What the previous call left behind is simply overwritten.
  return sglk::grouped_entry(A, B, sa, sb, C, expert_offsets, problem_sizes, num_experts, total_m, N, K, lda, ldb, ldc,
                             sa_stride_m, sa_stride_k, sb_stride_e, sb_stride_k, sb_stride_n, out_dtype, num_sms,
                             scratch, SGLK_GROUPED_STATIC_SCRATCH_BYTES, stream);
}

extern "C" int sglk_fp8_blockwise_scaled_mm_ex(const void* A, const void* B, const float* sa, const float* sb, void* C,
                                               int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                               int64_t sa_stride_m, int64_t sa_stride_k, int64_t sb_stride_k,
                                               int64_t sb_stride_n, int out_dtype, int variant, int num_sms, int flags,
                                               void* workspace, int64_t workspace_bytes, void* stream) {
  return sglk::gemm_entry(A, B, sa, sb, C, M, N, K, lda, ldb, ldc, sa_stride_m, sa_stride_k, sb_stride_k,
                          sb_stride_n, out_dtype, variant, num_sms, flags, workspace, workspace_bytes, stream);
}

extern "C" int sglk_fp8_blockwise_scaled_mm_allgather(const void* A, const void* B, const float* sa, const float* sb,
                                                      void* C_local, void* const* C_peers, int num_peers,
                                                      void* C_multicast, int64_t M,
                                                      int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                                      int64_t sa_stride_m, int64_t sa_stride_k, int64_t sb_stride_k,
                                                      int64_t sb_stride_n, int out_dtype, int variant, int num_sms,
                                                      void* workspace, int64_t workspace_bytes, void* stream) {
  if (variant < 0) variant = (M > 128 && N >= 256) ? 3 : 0;   // TMA-store variants only
  return sglk::gemm_entry(A, B, sa, sb, C_local, M, N, K, lda, ldb, ldc, sa_stride_m, sa_stride_k, sb_stride_k,
                          sb_stride_n, out_dtype, variant, num_sms, 0, workspace, workspace_bytes, stream,
                          C_multicast ? nullptr : C_peers, C_multicast ? 0 : num_peers, C_multicast);
}

extern "C" int sglk_fp8_blockwise_scaled_mm_quant(const void* A, const void* B, const float* sa, const float* sb,
                                                  void* out_q, float* out_s, int64_t M, int64_t N, int64_t K, int64_t lda,
                                                  int64_t ldb, int64_t ldq, int64_t sa_stride_m, int64_t sa_stride_k,
                                                  int64_t sb_stride_k, int64_t sb_stride_n, int64_t os_stride_m,
                                                  int64_t os_stride_g, int round_dtype, float eps, int silu_mul,
                                                  int num_sms, int flags, void* workspace, int64_t workspace_bytes,
                                                  void* stream) {
  sglk::QuantOut qo{out_q, ldq, out_s, os_stride_m, os_stride_g, eps, silu_mul ? 1 : 0};
  return sglk::gemm_entry(A, B, sa, sb, nullptr, M, N, K, lda, ldb, N, sa_stride_m, sa_stride_k, sb_stride_k, sb_stride_n,
                          round_dtype, -1, num_sms, flags, workspace, workspace_bytes, stream, nullptr, 0, nullptr, &qo);
}

extern "C" int sglk_fp8_blockwise_scaled_mm_gated(const void* A, const void* B, const float* sa, const float* sb, void* C,
                                                  int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                                  int64_t sa_stride_m, int64_t sa_stride_k, int64_t sb_stride_k,
                                                  int64_t sb_stride_n, int out_dtype, int num_sms,
                                                  const uint32_t* row_ready_counters, uint32_t ready_target,
                                                  int panel_rank, int panel_world, int whole_panels, int64_t wait_polls,
                                                  void* stream) {
  if (wait_polls < 0) return SGLK_E_ARG;
  const uint32_t polls = wait_polls > int64_t(0xFFFFFFFFu) ? 0xFFFFFFFFu : uint32_t(wait_polls);
  const sglk::GateIn gate{row_ready_counters, ready_target, panel_rank, panel_world, whole_panels, polls};
  return sglk::gemm_entry(A, B, sa, sb, C, M, N, K, lda, ldb, ldc, sa_stride_m, sa_stride_k, sb_stride_k, sb_stride_n,
                          out_dtype, 29, num_sms, 0, nullptr, 0, stream, nullptr, 0, nullptr, nullptr, &gate);
}

extern "C" int64_t sglk_gated_walk_order(int32_t* out, int64_t M, int64_t N, int64_t K, int panel_rank, int panel_world,
                                         int whole_panels) {
  using namespace sglk;
  if (M <= 0 || N < 256 || K <= 0 || panel_world < 1 || panel_rank < 0 || panel_rank >= panel_world ||
      M % (int64_t(2 * BLOCK_M) * panel_world) != 0) return -1;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.num_m_blocks = int32_t(M / (2 * BLOCK_M));
  p.num_n_blocks = int32_t((N + 255) / 256);
  int64_t gm = (int64_t(32) << 20) / (int64_t(2 * BLOCK_M) * K);        // as gemm_entry
  p.group_m = int32_t(gm < 4 ? 4 : (gm > 16 ? 16 : gm));
  set_panel_walk(p, panel_rank, panel_world, whole_panels);
  const int64_t tiles = int64_t(p.num_m_blocks) * p.num_n_blocks;
  if (out)
    for (int64_t t = 0; t < tiles; ++t) {
      int mb, nb;
      tile_coords_panels(int(t), p, mb, nb);
      out[2 * t] = mb;
      out[2 * t + 1] = nb;
    }
  return tiles;
}

extern "C" int64_t sglk_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int variant, int num_sms, int flags) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if (variant < 0) variant = sglk::pick_variant(M, N, K, true, num_sms);
  if (!sglk::variant_available(variant)) return 0;
  if (num_sms <= 0) num_sms = sglk::device_sm_count();
  if (variant == sglk::kRowsVariant || variant == sglk::kTRowsVariant) return 0;   // no split-K through global memory, no workspace
  if (variant == sglk::kDecodeVariant1 || variant == sglk::kDecodeVariant2)
    return sglk::decode_workspace_bytes(M, N, K, variant == sglk::kDecodeVariant2 ? 2 : 1, num_sms, flags);
  const sglk::Plan pl = sglk::make_plan(M, N, K, variant, num_sms, true, flags);
  const int64_t base = pl.stream_k ? pl.counter_bytes + pl.partial_bytes : (pl.scale_bytes ? sglk::kCounterAreaBytes : 0);
  return base + pl.scale_bytes;
}

extern "C" int sglk_fp8_blockwise_scaled_mm(const void* A, const void* B, const float* sa, const float* sb, void* C,
                                            int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                            int64_t sa_stride_m, int64_t sa_stride_k, int64_t sb_stride_k,
                                            int64_t sb_stride_n, int out_dtype, void* stream) {
  return sglk::gemm_entry(A, B, sa, sb, C, M, N, K, lda, ldb, ldc, sa_stride_m, sa_stride_k, sb_stride_k,
                          sb_stride_n, out_dtype, -1, 0, 0, nullptr, 0, stream);
}

extern "C" int sglk_gemm_num_variants(void) { return sglk::kNumVariants; }
extern "C" const char* sglk_gemm_variant_name(int v) {
  return (v >= 0 && v < sglk::kNumVariants) ? sglk::kVariants[v].name : "";
}
extern "C" int sglk_gemm_variant_is_product(int v) {
  return (v >= 0 && v < sglk::kNumVariants && sglk::kVariants[v].product) ? 1 : 0;
}
extern "C" int sglk_gemm_pick_variant(int64_t M, int64_t N, int64_t K) { return sglk::pick_variant(M, N, K, true); }
