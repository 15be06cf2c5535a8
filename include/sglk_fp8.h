/*
 * sglk_fp8.h — C ABI of libsglk_sm100a.so: the B200 (sm_100a) implementation of the FP8
 * block-quantized linear hot path of sgl-kernel.
 *
 * This is the drop-in boundary.  Plain pointers and sizes only — no torch types — so any host
 * (the Python `sgl_kernel` package in this repo via ctypes, or a C++ TORCH_LIBRARY shim, see
 * INTEGRATION.md) can bind it.  The shape of the interface follows the reference's own
 * "thin C-ABI" pattern for JIT kernels:
 *   /root/reference/include/sgl_kernel/jit_kernel/gemm/per_token_group_quant_8bit.hpp:272-302
 * (extern "C" launcher taking a queue pointer, raw data pointers and scalars), with a CUDA
 * stream in place of the SYCL queue.
 *
 * Conventions
 *   - All data pointers are DEVICE pointers (or pointers valid on the device).  Nothing is copied.
 *   - Work is enqueued on `stream` (a cudaStream_t passed as void*); no host synchronisation, no
 *     allocation, CUDA-graph capturable.  (Reference: "no .wait() -- PyTorch owns stream
 *     synchronization", ...8bit.hpp:265; src/sycl/Utils.h:75-77.)
 *   - Every function returns 0 on success, a negative SGLK_E_* code for a rejected argument
 *     (validated BEFORE any launch, like the reference's TORCH_CHECKs,
 *     src/sycl/per_token_group_quant_8bit.cpp:231-237,265-274,356) or a positive cudaError_t for a
 *     launch failure.  Nothing throws across the ABI.
 *   - There is no CPU fallback: on a machine without an sm_100 device the launch fails loudly.
 */
#ifndef SGLK_FP8_H_
#define SGLK_FP8_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SGLK_ABI_VERSION 3

#if defined(__GNUC__)
#define SGLK_API __attribute__((visibility("default")))
#else
#define SGLK_API
#endif

/* element types (in_dtype / out_dtype arguments) */
enum {
  SGLK_BF16 = 0,
  SGLK_F16 = 1,
  SGLK_F32 = 2,
  SGLK_E4M3 = 3, /* float8_e4m3fn */
  SGLK_INT8 = 4
};

/* layouts of output_s for the quant op (what the reference derives from output_s's dtype and
 * strides, src/sycl/per_token_group_quant_8bit.cpp:259-262,279-335) */
enum {
  SGLK_SCALE_F32_ROW_MAJOR = 0,  /* fp32 [M, G] contiguous                               */
  SGLK_SCALE_F32_COL_MAJOR = 1,  /* fp32, element (m,g) at g*scale_stride + m            */
  SGLK_SCALE_UE8M0_ROW_MAJOR = 2, /* uint8 exponent+127, [M, G] contiguous               */
  SGLK_SCALE_UE8M0_COL_PACKED = 3 /* 4 exponent bytes per int32, int32 (m, g/4) at (g/4)*scale_stride + m */
};

/* validation errors */
enum {
  SGLK_OK = 0,
  SGLK_E_NULL = -1,        /* null pointer                                   */
  SGLK_E_DTYPE = -2,       /* unsupported dtype                              */
  SGLK_E_GROUP = -3,       /* unsupported group size                         */
  SGLK_E_SHAPE = -4,       /* inconsistent / unsupported shape               */
  SGLK_E_STRIDE = -5,      /* stride or alignment not supported              */
  SGLK_E_LAYOUT = -6,      /* unknown scale layout                           */
  SGLK_E_ARG = -7,         /* other invalid argument                         */
  SGLK_E_DRIVER = -8,      /* cuTensorMapEncodeTiled unavailable / failed    */
  SGLK_E_VARIANT = -9,     /* unknown kernel variant                         */
  SGLK_E_CAPACITY = -10    /* problem larger than a fixed built-in buffer    */
};

SGLK_API int sglk_abi_version(void);
SGLK_API const char* sglk_error_string(int code);

/*
 * Replaces: at::native::xpu::sgl_per_token_group_quant_8bit
 *   /root/reference/src/sycl/per_token_group_quant_8bit.cpp:222-386  (host launcher)
 *   /root/reference/src/sycl/per_token_group_quant_8bit.cpp:48-220   (kernel)
 * and its JIT twin per_token_group_quant_8bit_forward_<in>_<out>_<gs>
 *   /root/reference/include/sgl_kernel/jit_kernel/gemm/per_token_group_quant_8bit.hpp:272-302.
 *
 * input     : [num_groups * group_size] contiguous elements of in_dtype (bf16 / f16 / f32)
 * output_q  : same number of 1-byte codes (e4m3fn or int8), contiguous
 * output_s  : scales, layout given by scale_layout; scale_stride in elements of the scale
 *             tensor's dtype (floats, or int32 words for SGLK_SCALE_UE8M0_COL_PACKED)
 * groups_per_row = hidden_dim / group_size (needed for the column-major layouts)
 * group_size in {16, 32, 64, 128, 256, 512}
 * Math: absmax = max(eps, max|x|); s = absmax / max_8bit (correctly rounded);
 *       UE8M0: s = 2^ceil(log2(max(s,1e-10))); q = clamp(x / s, min_8bit, max_8bit) -> RNE e4m3 /
 *       truncating int8.  Division is IEEE-correct (bit-exact with the torch oracle
 *       tests/test_per_token_group_quant_8bit.py:23-40, unlike the reference's x * (1/s)).
 */
SGLK_API int sglk_per_token_group_quant_8bit(const void* input, void* output_q, void* output_s,
                                    int64_t num_groups, int64_t groups_per_row, int64_t group_size,
                                    int in_dtype, int out_dtype, float eps, float min_8bit,
                                    float max_8bit, int scale_ue8m0, int scale_layout,
                                    int64_t scale_stride, void* stream);

/*
 * Replaces: at::native::xpu::sgl_per_token_group_quant_8bit_v2 for its fused-activation mode
 *   /root/reference/src/sycl/per_token_group_quant_8bit_v2.cpp:714-843 (host), :255-262 (silu*mul).
 * input is [M, 2*hidden] (gate | up); the quantised tensor is silu(gate)*up rounded to in_dtype,
 * [M, hidden].  Other arguments as above (groups are counted on the OUTPUT).
 */
SGLK_API int sglk_silu_mul_per_token_group_quant_8bit(const void* input, void* output_q, void* output_s,
                                             int64_t num_tokens, int64_t hidden, int64_t group_size,
                                             int in_dtype, int out_dtype, float eps, float min_8bit,
                                             float max_8bit, int scale_ue8m0, int scale_layout,
                                             int64_t scale_stride, void* stream);

/*
 * Replaces: at::native::xpu::sgl_per_token_group_quant_8bit_v2, whole surface
 *   /root/reference/src/sycl/per_token_group_quant_8bit_v2.cpp:714-843 (host), :176-330 (compute),
 *   :464-563 (masked MoE layout), schema src/torch_extension_sycl.cc:399-402.
 * input    : [num_experts, tokens_per_expert, hidden] (hidden*2 = gate|up when fuse_silu_and_mul), contiguous;
 *            num_experts == 1 unless masked_m is given
 * output_q : [num_experts, tokens_per_expert, hidden] codes
 * output_s : row-major layouts as in the v1 entry point.  Column-major layouts use the v2 op's addressing: the scale
 *            of (token t, group g) is written where the LOGICAL element (flat / G, flat % G), flat = t + g*tokens_per_expert,
 *            of the column-major tensor lives ((flat % G) [/4] * scale_stride + flat / G, + e*scale_expert_stride) —
 *            the "interleaved" pattern the reference's tests de-interleave
 *            (tests/test_per_token_group_quant_8bit_v2.py:807-817); packed UE8M0 words also get their unused tail
 *            bytes zeroed when G % 4 != 0 (…_v2.cpp:231-248).
 * masked_m : optional device int32[num_experts]; tokens >= masked_m[e] of expert e are skipped (outputs untouched).
 *            Only with column-major packed UE8M0 scales and fuse_silu_and_mul, like the reference (…_v2.cpp:779-783).
 * scale_expert_stride : elements of the scale tensor's dtype between expert slabs (output_s.stride(0)).
 */
SGLK_API int sglk_per_token_group_quant_8bit_v2(const void* input, void* output_q, void* output_s,
                                       int64_t num_experts, int64_t tokens_per_expert, int64_t hidden,
                                       int64_t group_size, int in_dtype, int out_dtype, float eps, float min_8bit,
                                       float max_8bit, int scale_ue8m0, int scale_layout, int64_t scale_stride,
                                       int64_t scale_expert_stride, int fuse_silu_and_mul, const int32_t* masked_m,
                                       void* stream);

/*
 * Replaces: at::native::xpu::sgl_per_token_quant_fp8(input, output_q, output_s)
 *   /root/reference/src/sycl/per_token_quant_fp8.cpp:201-306 (host), :69-199 (kernels); schema
 *   src/torch_extension_sycl.cc:410-411; python/sgl_kernel/gemm.py:236-241.
 * input    : [num_tokens, hidden_dim] contiguous bf16 / f16 (hidden_dim % 4 == 0, rows 8-byte aligned)
 * output_q : [num_tokens, hidden_dim] e4m3fn codes, contiguous
 * output_s : fp32 [num_tokens]: scale = rowmax|x| / 448
 * Math (reference order, :105-123): inv = scale == 0 ? 0 : 1/scale; q = e4m3_rne(clamp(float(x)*inv, +-448)) —
 * bit-exact with the reference's torch check (tests/test_per_token_quant_fp8.py:17-25).
 */
SGLK_API int sglk_per_token_quant_fp8(const void* input, void* output_q, void* output_s, int64_t num_tokens,
                             int64_t hidden_dim, int in_dtype, void* stream);

/*
 * Replaces: torch::Tensor fp8_blockwise_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype)
 *   declared /root/reference/include/sgl_kernel_ops.h:581-586, wrapped python/sgl_kernel/gemm.py:24-31,
 *   semantics pinned by tests/test_fp8_blockwise_gemm.py:23-63 (the reference ships no kernel).
 *
 * C[m,n] = sum_kb sa[m,kb] * sb[kb, n/128] * sum_{k in kb} A[m,k] * B[k,n]      (fp32 accumulate)
 * A  : e4m3fn [M,K], row-major, row stride lda (elements)
 * B  : e4m3fn [K,N] COLUMN-major == [N,K] row-major, stride ldb between consecutive n
 * sa : fp32 (m,kb) at m*sa_stride_m + kb*sa_stride_k   (M-major from the test, or quant's row-major)
 * sb : fp32 (kb,nb) at kb*sb_stride_k + nb*sb_stride_n, nb = n/128, ceil(N/128) columns
 * C  : bf16 or f16 [M,N], row stride ldc
 * Requirements: K % 16 == 0, lda % 16 == 0, ldb % 16 == 0, A/B 16-byte aligned.  Any M >= 0, N >= 0.
 */
SGLK_API int sglk_fp8_blockwise_scaled_mm(const void* A, const void* B, const float* sa, const float* sb,
                                 void* C, int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb,
                                 int64_t ldc, int64_t sa_stride_m, int64_t sa_stride_k,
                                 int64_t sb_stride_k, int64_t sb_stride_n, int out_dtype,
                                 void* stream);

/* flags of the _ex / _quant entry points and of sglk_gemm_workspace_bytes (0 = defaults).  They change scheduling only,
 * never which result is computed; there is no process-wide state in the library. */
enum {
  SGLK_GEMM_STREAMK_AUTO = 0,   /* cost model decides whether tiles are cut across CTAs (needs a workspace)          */
  SGLK_GEMM_STREAMK_FORCE = 1,  /* cut whenever the shape allows it (tests exercise the fix-up on small shapes)      */
  SGLK_GEMM_STREAMK_NEVER = 2,  /* whole tiles only                                                                 */
  SGLK_GEMM_GROUP_M_SHIFT = 8,  /* bits 8..15: m-blocks per rasterisation group of the prefill kernel, 0 = built-in */
  SGLK_GEMM_SPLIT_SHIFT = 16,   /* bits 16..23: decode kernels, how K is split when there are fewer tiles than SMs — workers per tile
                                   (swap-AB tile kernels) / CTAs per cluster, 1, 2 or 4 (row-balanced kernel); 0 = built-in */
  SGLK_GEMM_DECODE_2D_LOADS = 1 << 24   /* row-balanced decode kernel: one TMA load per K block even where one per stage would do (A/B hook) */
};

/* Same, with an explicit kernel variant (tile shape / CTA pairing) for tests and tuning, and an
 * optional device workspace that enables stream-K scheduling (decode shapes, ragged last waves):
 * variant = -1 selects the heuristic used by sglk_fp8_blockwise_scaled_mm; ids for which
 * sglk_gemm_variant_is_product() is 0 are laboratory variants that exist only in the diagnostics
 * build (libsglk_diag.so, -DSGLK_DIAG) — the product library answers SGLK_E_VARIANT for them;
 * num_sms <= 0 uses every SM of the current device; flags: SGLK_GEMM_* above;
 * workspace: >= sglk_gemm_workspace_bytes(...) bytes, 16-byte aligned, whose first MiB (the tile
 * counters) is ZERO on entry — the kernel leaves it zero, so one zero-initialised buffer can be
 * reused by consecutive launches (of any shape) on the same stream.  NULL disables stream-K. */
SGLK_API int sglk_fp8_blockwise_scaled_mm_ex(const void* A, const void* B, const float* sa, const float* sb,
                                    void* C, int64_t M, int64_t N, int64_t K, int64_t lda,
                                    int64_t ldb, int64_t ldc, int64_t sa_stride_m,
                                    int64_t sa_stride_k, int64_t sb_stride_k, int64_t sb_stride_n,
                                    int out_dtype, int variant, int num_sms, int flags, void* workspace,
                                    int64_t workspace_bytes, void* stream);

/*
 * GEMM with the NEXT layer's activation quantisation fused into its epilogue ("emit e4m3 + scales from
 * the previous GEMM's epilogue"; no reference counterpart — the reference runs fp8_blockwise_scaled_mm and
 * sgl_per_token_group_quant_8bit[_v2] as separate ops, python/sgl_kernel/gemm.py:24-31,85-126).
 * Result: bit-identical to  C = fp8_blockwise_scaled_mm(..., round_dtype);  [C = silu(C[:, gate]) * C[:, up];]
 *                           sgl_per_token_group_quant_8bit(C, out_q, out_s, 128, eps, -448, 448)
 * without C ever reaching HBM: each epilogue thread owns one row x 128 columns = one whole group.
 * silu_mul = 0: out_q [M, N] codes, out_s (row, g) for g < N/128.                        Needs N % 128 == 0.
 * silu_mul = 1: B's rows are interleaved by 128-row blocks, gate block j at rows [256j, 256j+128), up block j at
 *               [256j+128, 256j+256) (scales_b columns likewise); out_q [M, N/2], out_s (row, g) for g < N/256.
 *               silu in fp32 rounded to round_dtype, product rounded to round_dtype — the v2 quant op's fused mode
 *               (/root/reference/src/sycl/per_token_group_quant_8bit_v2.cpp:255-262).            Needs N % 256 == 0.
 *               LABORATORY BUILD ONLY (libsglk_diag.so): correct but slower than the separate ops (the exact expf +
 *               division chain runs in the epilogue while the tensor pipe waits); the product library answers
 *               SGLK_E_VARIANT.
 * out_q row stride ldq (bytes, multiple of 16); out_s element (row, g) at row*os_stride_m + g*os_stride_g floats.
 * Other arguments as sglk_fp8_blockwise_scaled_mm_ex.
 */
SGLK_API int sglk_fp8_blockwise_scaled_mm_quant(const void* A, const void* B, const float* sa, const float* sb,
                                       void* out_q, float* out_s, int64_t M, int64_t N, int64_t K, int64_t lda,
                                       int64_t ldb, int64_t ldq, int64_t sa_stride_m, int64_t sa_stride_k,
                                       int64_t sb_stride_k, int64_t sb_stride_n, int64_t os_stride_m,
                                       int64_t os_stride_g, int round_dtype, float eps, int silu_mul, int num_sms,
                                       int flags, void* workspace, int64_t workspace_bytes, void* stream);

/*
 * Fused GEMM -> all-gather for the column-(N-)sharded linear (no reference counterpart: the reference
 * has no communication layer).  Same arithmetic as above on this rank's [K, N_local]
 * weight panel; the epilogue TMA-stores every finished 32x64 output box to C_local AND to the same
 * position of each peer's buffer, so the [M, N_total] result is assembled on all GPUs while the MMAs
 * still run.  C_local / C_peers[i] point at THIS rank's panel inside each GPU's row-major [M, N_total]
 * buffer (base + panel_offset), ldc = N_total; peers' pointers must be peer-mapped into this process
 * (CUDA VMM / symmetric memory / IPC) and 16-byte aligned.  The caller synchronises the ranks before
 * the buffers are read (the kernel guarantees only that its own stores have completed at kernel end).
 * num_peers <= 7.
 * C_multicast (optional): the NVLS multicast address of the same panel position (a multicast mapping of
 * the symmetric buffer that includes this GPU).  When non-NULL the epilogue issues multimem.st to it
 * instead of per-peer stores — each output byte leaves this GPU once and the NVSwitch replicates it —
 * and C_local / C_peers are not written separately.  Needs N % 8 == 0.
 */
SGLK_API int sglk_fp8_blockwise_scaled_mm_allgather(const void* A, const void* B, const float* sa, const float* sb,
                                           void* C_local, void* const* C_peers, int num_peers,
                                           void* C_multicast, int64_t M,
                                           int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                           int64_t sa_stride_m, int64_t sa_stride_k, int64_t sb_stride_k,
                                           int64_t sb_stride_n, int out_dtype, int variant, int num_sms,
                                           void* workspace, int64_t workspace_bytes, void* stream);

/*
 * Row-sharded activation quantisation for the N-sharded layer on several GPUs, and the GEMM that consumes it while it
 * arrives (no reference counterpart: the reference has no communication layer).  Replicated activations
 * make every rank quantise the whole batch; here rank r quantises only ITS rows — the same arithmetic as
 * sglk_per_token_group_quant_8bit (group 128, e4m3, row-major fp32 scales; bit-identical, quantisation is row-local) —
 * and the codes and scales land in EVERY rank's copy of a symmetric [M, hidden] / [M, hidden/128] buffer:
 *   input            this rank's rows, [rows, hidden] bf16/f16 contiguous, 32-byte aligned
 *   output_q/_s      where those rows live in the code / scale buffer (base + row offset); with multicast != 0 these
 *                    are NVLS multicast addresses (one multimem.st reaches every GPU), else plain device pointers
 *   ready_counters   the counter of the first of these rows' blocks (same addressing rule).  Every rows_per_counter rows
 *                    share one uint32 counter; it is incremented (release, system scope) once per finished chunk of 16384
 *                    elements, rows_per_counter * hidden / 16384 times per block and call.  Counters are never reset.
 * Requirements: hidden % 128 == 0, rows % rows_per_counter == 0, rows_per_counter * hidden % 16384 == 0.
 */
SGLK_API int sglk_per_token_group_quant_fp8_bcast(const void* input, void* output_q, float* output_s,
                                         uint32_t* ready_counters, int64_t rows, int64_t hidden,
                                         int64_t rows_per_counter, int in_dtype, float eps, int multicast,
                                         void* stream);

/*
 * sglk_fp8_blockwise_scaled_mm whose A rows (codes and row-major scales) ARRIVE WHILE THE KERNEL RUNS: one arrival
 * counter per 256 rows (row_ready_counters[m / 256], as written by sglk_per_token_group_quant_fp8_bcast with
 * rows_per_counter = 256); the 256-row tiles of a block are started once int32(counter - ready_target) >= 0, i.e. the
 * caller passes (calls so far on this buffer + 1) * 256 * K / 16384.  The blocks are walked in arrival order: the
 * panel of rank panel_rank first (rows [rank, rank + 1) * M / world), then either (whole_panels == 0: every rank
 * broadcasts at once) the j-th block of every other panel, nearest rank first, for j = 0, 1, ..., or (whole_panels != 0:
 * ring schedule, one sender per receiver at a time) the panels of ranks rank+1, rank+2, ... one after the other.
 * Same result bits as sglk_fp8_blockwise_scaled_mm on the completed buffers.
 * Requirements: M % (256 * panel_world) == 0, N >= 256, K-block-contiguous scales (sa_stride_k == sb_stride_k == 1,
 * 16-byte aligned rows), C 16-byte aligned rows.
 * wait_polls bounds how long a CTA waits for a counter: 0 = for ever (what a collective normally wants: a peer held up by
 * a module load, a host pause or an allocator call is merely late, and the launch finishes when it delivers — a peer
 * that is really gone has to be dealt with by the host's watchdog, as with NCCL); n > 0 = after n polling rounds (about
 * 1-2 us each: 4 000 000 is a few seconds) the CTA executes a trap, which fails the launch with a STICKY error — the
 * CUDA context of this process is unusable afterwards and has to be torn down.  Use a bound only where a dead context
 * is preferable to a hang (tests, CI boxes with a wall-clock limit).
 */
SGLK_API int sglk_fp8_blockwise_scaled_mm_gated(const void* A, const void* B, const float* sa, const float* sb, void* C,
                                       int64_t M, int64_t N, int64_t K, int64_t lda, int64_t ldb, int64_t ldc,
                                       int64_t sa_stride_m, int64_t sa_stride_k, int64_t sb_stride_k,
                                       int64_t sb_stride_n, int out_dtype, int num_sms,
                                       const uint32_t* row_ready_counters, uint32_t ready_target, int panel_rank,
                                       int panel_world, int whole_panels, int64_t wait_polls, void* stream);

/* Host-side copy of the gated kernel's walk, for tests and schedulers: out[2t] / out[2t+1] = the 256-row m-block and the
 * 256-column n-block of the t-th tile in launch order (tile t goes to CTA pair t mod #pairs).  Returns the number of
 * tiles (out may be NULL to ask for it), -1 for shapes the gated entry point rejects.  No device is touched. */
SGLK_API int64_t sglk_gated_walk_order(int32_t* out, int64_t M, int64_t N, int64_t K, int panel_rank, int panel_world,
                              int whole_panels);

/*
 * Replaces: void fp8_blockwise_scaled_grouped_mm(output, a_ptrs, ..., problem_sizes, expert_offsets, workspace)
 *   declared /root/reference/include/sgl_kernel_ops.h:660-678, wrapped python/sgl_kernel/moe.py:236-276, semantics
 *   pinned by tests/test_fp8_blockwise_moe.py:27-51,66-171 (the reference ships no kernel).
 * For every expert e: C[off_e : off_e + m_e, :] = blockwise_scaled_mm(A[off_e : off_e + m_e], B[e], sa[off_e : ...], sb[e]).
 * A  : e4m3 [total_m, K] — the experts' activations stacked, row stride lda
 * B  : e4m3 [num_experts * N, K] — expert e's [N, K] (K-contiguous) weight at rows [e*N, (e+1)*N), row stride ldb
 * sa : fp32 (row, kb) at row*sa_stride_m + kb*sa_stride_k (rows of the stack)
 * sb : fp32 (e, kb, nb) at e*sb_stride_e + kb*sb_stride_k + nb*sb_stride_n
 * C  : bf16 / f16 [total_m, N], row stride ldc
 * expert_offsets : DEVICE int32[num_experts], first stacked row of each expert
 * problem_sizes  : DEVICE int32[num_experts, 3] = (m_e, n, k); only m_e is read (n, k are the N, K arguments)
 * workspace : >= sglk_grouped_mm_workspace_bytes(num_experts, total_m) bytes, 16-byte aligned; the m-tile table is
 *             built there on the device — no host synchronisation, CUDA-graph capturable.
 */
SGLK_API int sglk_fp8_blockwise_scaled_grouped_mm(const void* A, const void* B, const float* sa, const float* sb, void* C,
                                         const int32_t* expert_offsets, const int32_t* problem_sizes,
                                         int64_t num_experts, int64_t total_m, int64_t N, int64_t K, int64_t lda,
                                         int64_t ldb, int64_t ldc, int64_t sa_stride_m, int64_t sa_stride_k,
                                         int64_t sb_stride_e, int64_t sb_stride_k, int64_t sb_stride_n, int out_dtype,
                                         int num_sms, void* workspace, int64_t workspace_bytes, void* stream);
SGLK_API int64_t sglk_grouped_mm_workspace_bytes(int64_t num_experts, int64_t total_m);

/*
 * The grouped GEMM above for callers that have no way to give it a workspace.  No counterpart in the reference (this is
 * an addition of this library, not a replacement for anything there).
 *
 * Who needs it: hosts that may not create device memory at the point of the call — a plugin that runs inside someone
 * else's memory pool and is forbidden to call cudaMalloc, a serving loop that has handed all of HBM to its KV cache and
 * must not fragment it, a thin FFI binding (cgo / JNI / ctypes) that receives input and output pointers from its
 * runtime but owns no allocator of its own.  The caller passes exactly the inputs and the output; the m-tile table the
 * kernel needs is built in an array that is part of the library image (SGLK_GROUPED_STATIC_SCRATCH_BYTES, on every GPU
 * that loads the library, from load until the process exits).  The function and everything under it call no cudaMalloc,
 * cudaFree or other allocating routine.
 *
 * Limit: ceil(total_m / 128) + num_experts <= sglk_grouped_mm_static_max_tiles() (4095 with the 64 KiB array: e.g. 256
 * experts and up to 491 392 stacked rows).  Anything larger is refused with SGLK_E_CAPACITY before any launch — use the
 * workspace form for it.
 *
 * First call on each device: the library asks the CUDA runtime where that device's copy of the array lives (once per
 * device, with that device current, i.e. the device of `stream`).  Make that first call outside any stream capture;
 * every later call on the device only reads the remembered address.
 *
 * Sharing: all calls on a device use the SAME array, and it keeps what the last call wrote.  Nothing in it has to be
 * zero on entry — each call rebuilds the tile count and every table entry it reads, in stream order, before its GEMM
 * kernel starts — so back-to-back calls on one stream are fine.  Two calls that can run on the GPU at the same time
 * (different streams without an event between them, or several host threads each on their own stream) would overwrite
 * each other's table: that is a race the library does not guard against.  Serialise such calls, or use the workspace
 * form with one workspace per stream.
 *
 * Arguments and result bits are those of sglk_fp8_blockwise_scaled_grouped_mm (same kernels, same tile order).
 */
#define SGLK_GROUPED_STATIC_SCRATCH_BYTES 65536
SGLK_API int sglk_fp8_blockwise_scaled_grouped_mm_static(const void* A, const void* B, const float* sa, const float* sb,
                                                void* C, const int32_t* expert_offsets, const int32_t* problem_sizes,
                                                int64_t num_experts, int64_t total_m, int64_t N, int64_t K, int64_t lda,
                                                int64_t ldb, int64_t ldc, int64_t sa_stride_m, int64_t sa_stride_k,
                                                int64_t sb_stride_e, int64_t sb_stride_k, int64_t sb_stride_n,
                                                int out_dtype, int num_sms, void* stream);
/* largest ceil(total_m / 128) + num_experts the call above accepts */
SGLK_API int64_t sglk_grouped_mm_static_max_tiles(void);

/* bytes of workspace the call above can make use of (0: none): [1 MiB tile counters][stream-K partials, if the
 * scheduler would cut tiles][K-block-contiguous copies of sa / sb, for prefill-sized problems whose scales are not
 * already laid out that way] */
SGLK_API int64_t sglk_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K, int variant, int num_sms, int flags);

/* number of GEMM variant ids, a printable name for each ("" if out of range), and whether the id is a product
 * variant (1) or a laboratory one that only the diagnostics build can launch (0) */
SGLK_API int sglk_gemm_num_variants(void);
SGLK_API const char* sglk_gemm_variant_name(int variant);
SGLK_API int sglk_gemm_variant_is_product(int variant);
/* the variant the heuristic would pick for a problem with a plain [M, N] output */
SGLK_API int sglk_gemm_pick_variant(int64_t M, int64_t N, int64_t K);

/* Device self-check of the quant kernel's division scheme against __fdiv_rn over every 16-bit
 * numerator (numerator_dtype = SGLK_BF16 or SGLK_F16) for `num_scales` scales; adds the number of
 * differing e4m3 codes to *mismatches (device ptr). */
SGLK_API int sglk_selftest_division(const float* scales, int64_t num_scales, int numerator_dtype,
                           unsigned long long* mismatches, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SGLK_FP8_H_ */
