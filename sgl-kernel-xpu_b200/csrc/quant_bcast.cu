// Row-sharded group quantisation for the N-sharded linear layer on several GPUs: every rank quantises ITS rows of the
// batch and the codes + scales land in every rank's copy of a symmetric [M, K] / [M, K/128] buffer, with a per-256-row
// arrival counter that the consuming GEMM polls (sglk_fp8_blockwise_scaled_mm_gated).  No reference counterpart (the
// reference has no communication layer); the arithmetic IS the group-quant kernel's — the same
// ptgq_chunk body (quant_kernel.cuh) instantiated with multicast stores — so the result is bit-identical to
// sgl_per_token_group_quant_8bit on the whole batch (/root/reference/src/sycl/per_token_group_quant_8bit.cpp:48-220).
//
//   * stores: multimem.st to the NVLS multicast mapping — each byte leaves the GPU once, the NVSwitch replicates it
//     (multicast == 0: plain stores to a local buffer — the single-GPU form the tests drive);
//   * signalling: one CTA = one chunk of 1024 16-element vectors; 256 rows are a whole number of chunks (hidden % 64
//     == 0), so after its stores a CTA adds 1 to the counter of its 256-row block on every GPU:
//     bar.sync -> thread 0: fence.acq_rel.sys -> multimem.red.release.sys.add.  Counters only grow; the consumer of step
//     t waits for (t + 1) * chunks-per-block — nothing is ever reset, so nothing races with a reset.
#include "quant_kernel.cuh"

namespace sglk {

template <typename T, int kSt>
__global__ void __launch_bounds__(256)
ptgq_bcast_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ s, const QuantParams p,
                  uint32_t* __restrict__ flags, int chunks_per_flag) {
  using C = QuantCfg<T, 128, false, SGLK_SCALE_F32_ROW_MAJOR, false>;
  const int64_t total_slots = p.num_groups * C::LPG;
  ptgq_chunk<T, 128, false, SGLK_SCALE_F32_ROW_MAJOR, false, false, true, kSt>(
      x, q, s, p, int64_t(blockIdx.x) * C::SLOTS_PER_CHUNK, total_slots, int(threadIdx.x));
  __syncthreads();                                   // every thread's stores happen-before thread 0's fence
  if (threadIdx.x == 0) {
    uint32_t* f = flags + blockIdx.x / unsigned(chunks_per_flag);
    asm volatile("fence.acq_rel.sys;" ::: "memory");
    if constexpr (kSt == 1) asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(f), "r"(1u) : "memory");
    else asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(f), "r"(1u) : "memory");
  }
}

template <typename T>
static int launch_bcast(const void* x, void* q, float* s, const QuantParams& p, uint32_t* flags, int chunks_per_flag,
                        int64_t chunks, int multicast, cudaStream_t st) {
  if (multicast)
    ptgq_bcast_kernel<T, 1><<<unsigned(chunks), 256, 0, st>>>(static_cast<const T*>(x), static_cast<uint8_t*>(q), s, p, flags,
                                                              chunks_per_flag);
  else
    ptgq_bcast_kernel<T, 0><<<unsigned(chunks), 256, 0, st>>>(static_cast<const T*>(x), static_cast<uint8_t*>(q), s, p, flags,
                                                              chunks_per_flag);
  return int(cudaGetLastError());
}

}  // namespace sglk

extern "C" int sglk_per_token_group_quant_fp8_bcast(const void* input, void* output_q, float* output_s,
                                                    uint32_t* ready_counters, int64_t rows, int64_t hidden,
                                                    int64_t rows_per_counter, int in_dtype, float eps, int multicast,
                                                    void* stream) {
  using namespace sglk;
  if (rows < 0 || hidden <= 0 || rows_per_counter <= 0) return SGLK_E_SHAPE;
  if (rows == 0) return SGLK_OK;
  if (!input || !output_q || !output_s || !ready_counters) return SGLK_E_NULL;
  if (in_dtype != SGLK_BF16 && in_dtype != SGLK_F16) return SGLK_E_DTYPE;
  // whole groups, whole counter blocks, and a counter block must be a whole number of 1024-vector chunks
  constexpr int64_t kChunkVectors = 1024;
  if (hidden % 128 != 0 || rows % rows_per_counter != 0 || (rows_per_counter * (hidden / 16)) % kChunkVectors != 0) return SGLK_E_SHAPE;
  if (rows * (hidden / 128) >= (int64_t(1) << 31)) return SGLK_E_SHAPE;
  if ((reinterpret_cast<uintptr_t>(input) & 31) || (reinterpret_cast<uintptr_t>(output_q) & 15) ||
      (reinterpret_cast<uintptr_t>(output_s) & 3) || (reinterpret_cast<uintptr_t>(ready_counters) & 3)) return SGLK_E_STRIDE;
  QuantParams p;
  memset(&p, 0, sizeof(p));
  p.num_groups = rows * (hidden / 128);
  p.groups_per_row = uint32_t(hidden / 128);
  p.in_row_stride = hidden;
  p.eps = eps; p.qmin = -448.0f; p.qmax = 448.0f;
  p.rcp_qmax = 1.0f / 448.0f;                       // correctly rounded on the host, as in quant_dispatch
  p.tokens_per_expert = uint32_t(rows);
  const int64_t chunks = rows * (hidden / 16) / kChunkVectors;
  const int chunks_per_flag = int(rows_per_counter * (hidden / 16) / kChunkVectors);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (in_dtype == SGLK_BF16)
    return launch_bcast<__nv_bfloat16>(input, output_q, output_s, p, ready_counters, chunks_per_flag, chunks, multicast, st);
  return launch_bcast<__half>(input, output_q, output_s, p, ready_counters, chunks_per_flag, chunks, multicast, st);
}
