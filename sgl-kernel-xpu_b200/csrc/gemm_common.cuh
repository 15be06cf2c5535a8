// Pieces shared by the prefill (gemm.cu) and decode (gemm_decode.cu) kernels: operand / instruction descriptors,
// output packing, tensor-map encoding, the stream-K workspace layout.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <mutex>

#include "../../include/sglk_fp8.h"
#include "common.cuh"
#include "ptx.cuh"

namespace sglk {

constexpr int BLOCK_M = 128;   // rows per CTA (TMEM lanes)
constexpr int BLOCK_K = 128;   // one scale block; 128 B of e4m3 = one 128B-swizzle atom row
constexpr int UMMA_K = 32;     // K per tcgen05.mma for 8-bit operands
constexpr int kMaxDynSmem = 232448;                    // 227 KB
constexpr int64_t kCounterAreaBytes = 1 << 20;   // first MiB of a stream-K workspace: tile counters (zero between launches)

// K-major, 128-byte-swizzled operand tile: 8-row x 128-B atoms, 1024 B apart (SBO); LBO unused.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= uint64_t((smem_addr & 0x3FFFFu) >> 4);        // start address      bits [0,14)
  d |= uint64_t(0) << 16;                            // leading byte offset bits [16,30) (ignored)
  d |= uint64_t(1024 >> 4) << 32;                    // stride byte offset  bits [32,46)
  d |= uint64_t(1) << 46;                            // descriptor version (sm_100)
  d |= uint64_t(2) << 61;                            // SWIZZLE_128B
  return d;
}
// kind::f8f6f4, A = B = e4m3, D = f32, both K-major, dense
__host__ __device__ constexpr uint32_t make_idesc(int umma_m, int umma_n) {
  return (1u << 4)                        // D format f32
       | (0u << 7) | (0u << 10)           // A, B format e4m3
       | (0u << 15) | (0u << 16)          // A, B K-major
       | (uint32_t(umma_n >> 3) << 17)
       | (uint32_t(umma_m >> 4) << 24);
}

template <typename OutT> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&v);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  const __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&v);
}


// --------------------------------------------------------------------------------------------
// host side
// --------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// 2-D row-major tensor [rows, cols] with `ld` elements between rows; box = [box_rows, box_cols]
inline bool make_tmap_2d(CUtensorMap* m, CUtensorMapDataType dt, int elem_bytes, const void* ptr, uint64_t rows,
                         uint64_t cols, uint64_t ld, uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return false;
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {ld * uint64_t(elem_bytes)};
  const cuuint32_t box[2] = {box_cols, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return enc(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// The same row-major [rows, cols] byte matrix (cols % 128 == 0) seen as [kblock][row][128 B]: one box of
// [box_kblocks][box_rows][128 B] fetches `box_kblocks` consecutive 128-byte K blocks of `box_rows` rows with ONE TMA
// instruction, and lands them as box_kblocks back-to-back K-major swizzled tiles (box_rows % 8 == 0 keeps every tile
// 1 KB aligned).  Coordinates: {0, row, kblock}.
inline bool make_tmap_kblocks(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
                              uint32_t box_kblocks) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc || cols % 128 != 0) return false;
  const cuuint64_t dims[3] = {128, rows, cols / 128};
  const cuuint64_t strides[2] = {ld, 128};
  const cuuint32_t box[3] = {128, box_rows, box_kblocks};
  const cuuint32_t estr[3] = {1, 1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(ptr), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}


// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of the current device's context: opt in once per
// (kernel instantiation, device ordinal).  `configured` is the instantiation's own flag word.
template <typename Kern>
inline int ensure_max_dyn_smem(Kern kern, std::atomic<uint64_t>& configured, int smem) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = -1;
  if (dev < 0 || dev >= 64 || !((configured.load(std::memory_order_acquire) >> dev) & 1)) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return int(e);
    if (dev >= 0 && dev < 64) configured.fetch_or(uint64_t(1) << dev, std::memory_order_release);
  }
  return 0;
}

// decode path (gemm_decode.cu): swap-AB kernel, weights on the MMA M side.  Same argument meaning as gemm_entry.
int decode_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M, int64_t N,
                 int64_t K, int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk, int64_t sb_sk,
                 int64_t sb_sn, int out_dtype, int cg, int num_sms, int flags, void* workspace, int64_t workspace_bytes,
                 cudaStream_t st);
int64_t decode_workspace_bytes(int64_t M, int64_t N, int64_t K, int cg, int num_sms, int flags);
bool decode_supported(int64_t M, int64_t N, int64_t K);
// row-balanced swap-AB kernel (gemm_decode.cu): every CTA owns one contiguous range of weight rows for all of K
int rows_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M, int64_t N, int64_t K,
               int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk, int64_t sb_sk, int64_t sb_sn,
               int out_dtype, int num_sms, int flags, cudaStream_t st);
bool rows_supported(int64_t M, int64_t N, int64_t K);
// row-balanced token-major kernel (gemm_trows.cu): the same split with the tokens on the MMA's M side, for 65..128 tokens
int trows_entry(const void* A, const void* B, const float* sa, const float* sb, void* Cout, int64_t M, int64_t N, int64_t K,
                int64_t lda, int64_t ldb, int64_t ldc, int64_t sa_sm, int64_t sa_sk, int64_t sb_sk, int64_t sb_sn,
                int out_dtype, int num_sms, cudaStream_t st);
bool trows_supported(int64_t M, int64_t N, int64_t K);

}  // namespace sglk
