"""torch.ops.sgl_kernel.* registrations for the FP8 block-quantized linear path (dispatch key CUDA).

Schemas are the reference's, verbatim:
  sgl_per_token_group_quant_8bit / _v2 : /root/reference/src/torch_extension_sycl.cc:395-402
  fp8_blockwise_scaled_mm              : derived from include/sgl_kernel_ops.h:581-586 +
                                         python/sgl_kernel/gemm.py:24-31 (the reference registers none)
The reference binds dispatch key XPU; this build binds CUDA.  Each impl validates like the reference's
host launcher (TORCH_CHECK -> RuntimeError before any launch), then calls the C ABI with raw pointers
and the *current* CUDA stream.  No host sync, no allocation in quant, caching-allocator only in GEMM.
"""
from __future__ import annotations

import threading
from typing import Optional

import torch

from . import _lib

_IN_DTYPES = {torch.bfloat16: _lib.BF16, torch.float16: _lib.F16, torch.float32: _lib.F32}
_OUT_DTYPES = {torch.float8_e4m3fn: _lib.E4M3, torch.int8: _lib.INT8}
_SUPPORTED_GROUPS_V1 = (32, 64, 128, 256, 512)   # per_token_group_quant_8bit.cpp:337-358
_SUPPORTED_GROUPS_V2 = (16, 32, 64, 128)         # per_token_group_quant_8bit_v2.cpp:829-839

_QUANT_SCHEMA = (
    "sgl_per_token_group_quant_8bit(Tensor input, Tensor output_q, Tensor output_s, int group_size,"
    " float eps, float fp8_min, float fp8_max, bool scale_ue8m0) -> ()"
)
_QUANT_V2_SCHEMA = (
    "sgl_per_token_group_quant_8bit_v2(Tensor input, Tensor output_q, Tensor output_s, int group_size,"
    " float eps, float fp8_min, float fp8_max, bool scale_ue8m0, bool fuse_silu_and_mul, Tensor? masked_m) -> ()"
)
_GEMM_SCHEMA = (
    "fp8_blockwise_scaled_mm(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b,"
    " ScalarType out_dtype) -> Tensor"
)

# derived from the C++ declaration include/sgl_kernel_ops.h:660-678 + python/sgl_kernel/moe.py:236-276 (the reference
# registers none); `output` is written in place, the pointer / stride / layout tensors are the CUTLASS scratch of the
# upstream CUDA op and are accepted but not used here
_GROUPED_SCHEMA = (
    "fp8_blockwise_scaled_grouped_mm(Tensor(a!) output, Tensor a_ptrs, Tensor b_ptrs, Tensor out_ptrs,"
    " Tensor a_scales_ptrs, Tensor b_scales_ptrs, Tensor a, Tensor b, Tensor scales_a, Tensor scales_b,"
    " Tensor stride_a, Tensor stride_b, Tensor stride_c, Tensor layout_sfa, Tensor layout_sfb,"
    " Tensor problem_sizes, Tensor expert_offsets, Tensor workspace) -> ()"
)

# verbatim: src/torch_extension_sycl.cc:410
_TOKEN_QUANT_SCHEMA = "sgl_per_token_quant_fp8(Tensor input, Tensor(a!) output_q, Tensor(b!) output_s) -> ()"

# B200 extension (no reference counterpart): the GEMM with the next layer's group quantisation fused into its epilogue
_GEMM_QUANT_SCHEMA = (
    "fp8_blockwise_scaled_mm_quant(Tensor mat_a, Tensor mat_b, Tensor scales_a, Tensor scales_b, ScalarType round_dtype,"
    " float eps, bool fuse_silu_and_mul, bool column_major_scales) -> (Tensor, Tensor)"
)

_library = torch.library.Library("sgl_kernel", "FRAGMENT")
_library.define(_GEMM_QUANT_SCHEMA)
_library.define(_TOKEN_QUANT_SCHEMA)
_library.define(_QUANT_SCHEMA)
_library.define(_QUANT_V2_SCHEMA)
_library.define(_GEMM_SCHEMA)
_library.define(_GROUPED_SCHEMA)


def _stream_ptr(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _check(cond: bool, msg: str) -> None:
    if not cond:
        raise RuntimeError(msg)


def _scale_layout(output_s: torch.Tensor, scale_ue8m0: bool):
    """What the reference derives from output_s (per_token_group_quant_8bit.cpp:237,259-262,279-335)."""
    _check(output_s.dim() == 2, "Expected output_s.dim() == 2")
    col_major = output_s.stride(0) < output_s.stride(1)
    stride = output_s.stride(1)
    if scale_ue8m0:
        if col_major:
            _check(output_s.dtype == torch.int32, "UE8M0 column-major scales must be packed int32")
            return _lib.SCALE_UE8M0_COL_PACKED, stride
        _check(output_s.dtype in (torch.uint8, getattr(torch, "float8_e8m0fnu", torch.uint8)),
               "UE8M0 row-major scales must be uint8")
        _check(output_s.is_contiguous(), "UE8M0 row-major scales must be contiguous")
        return _lib.SCALE_UE8M0_ROW_MAJOR, stride
    _check(output_s.dtype == torch.float32, "output_s must be float32 unless scale_ue8m0")
    if col_major:
        return _lib.SCALE_F32_COL_MAJOR, stride
    _check(output_s.is_contiguous(), "row-major output_s must be contiguous")
    return _lib.SCALE_F32_ROW_MAJOR, stride


def _quant_common(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0, fuse, groups_ok, name):
    _check(input.is_cuda and output_q.is_cuda and output_s.is_cuda, f"{name}: tensors must be CUDA tensors")
    _check(input.is_contiguous(), "input must be contiguous")          # CHECK_CONTIGUOUS, Utils.h:13
    _check(output_q.is_contiguous(), "output_q must be contiguous")
    _check(input.dtype in _IN_DTYPES,
           f"{name}: input dtype must be Float16, BFloat16, or Float32, got {input.dtype}")
    _check(output_q.dtype in _OUT_DTYPES,
           f"{name}: output_q dtype must be Int8 or Float8_e4m3fn, got {output_q.dtype}")
    _check(group_size in groups_ok, f"Unsupported group_size: {group_size}. Supported sizes are: {groups_ok}")
    hidden_in = input.size(-1) if input.dim() > 0 else 1
    hidden = hidden_in // 2 if fuse else hidden_in
    _check(not fuse or hidden_in % 2 == 0, "fuse_silu_and_mul needs an even hidden size")
    _check(hidden % group_size == 0 and input.numel() % group_size == 0,
           "input.numel() must be divisible by group_size")
    _check(output_q.numel() == (input.numel() // 2 if fuse else input.numel()),
           "output_q must have one code per quantised element")
    layout, stride = _scale_layout(output_s, scale_ue8m0)
    num_tokens = input.numel() // hidden_in if hidden_in else 0
    groups_per_row = hidden // group_size
    if layout in (_lib.SCALE_F32_ROW_MAJOR, _lib.SCALE_UE8M0_ROW_MAJOR):
        _check(output_s.numel() >= num_tokens * groups_per_row, "output_s too small")
    else:
        _check(output_s.size(0) >= num_tokens, "output_s has fewer rows than tokens")
    if input.numel() == 0:
        return
    lib = _lib.load()
    with torch.cuda.device(input.device):
        if fuse:
            rc = lib.sglk_silu_mul_per_token_group_quant_8bit(
                input.data_ptr(), output_q.data_ptr(), output_s.data_ptr(), num_tokens, hidden, group_size,
                _IN_DTYPES[input.dtype], _OUT_DTYPES[output_q.dtype], eps, fp8_min, fp8_max,
                int(scale_ue8m0), layout, stride, _stream_ptr(input))
        else:
            rc = lib.sglk_per_token_group_quant_8bit(
                input.data_ptr(), output_q.data_ptr(), output_s.data_ptr(), num_tokens * groups_per_row,
                groups_per_row, group_size, _IN_DTYPES[input.dtype], _OUT_DTYPES[output_q.dtype], eps,
                fp8_min, fp8_max, int(scale_ue8m0), layout, stride, _stream_ptr(input))
    _lib.check(rc, name)


def _quant_v1(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0):
    """at::native::xpu::sgl_per_token_group_quant_8bit, src/sycl/per_token_group_quant_8bit.cpp:222-386."""
    _quant_common(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0, False,
                  _SUPPORTED_GROUPS_V1 + (16,), "sgl_per_token_group_quant_8bit")


def _quant_v2(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0,
              fuse_silu_and_mul, masked_m: Optional[torch.Tensor]):
    """sgl_per_token_group_quant_8bit_v2, src/sycl/per_token_group_quant_8bit_v2.cpp:714-843.

    Differences from the v1 op that a caller can see, all reproduced: eps must be 1e-10 (:732) and (min, max) the
    destination type's own range (:780-781); column-major scales are written with the op's interleaved addressing
    (:200-229); `masked_m` selects the masked MoE layout — input [E, T, 2*hidden], 3-D output_s, only tokens
    < masked_m[e] quantised — which the reference dispatches only for column-major packed UE8M0 scales together
    with fuse_silu_and_mul (:783-790)."""
    name = "sgl_per_token_group_quant_8bit_v2"
    fuse = bool(fuse_silu_and_mul)
    _check(abs(1e-10 - eps) < 1e-13, f"{name}: eps must be 1e-10")
    _check(input.is_cuda and output_q.is_cuda and output_s.is_cuda, f"{name}: tensors must be CUDA tensors")
    _check(output_q.dtype in _OUT_DTYPES, f"{name}: output_q dtype must be Int8 or Float8_e4m3fn, got {output_q.dtype}")
    lo, hi = (-448.0, 448.0) if output_q.dtype == torch.float8_e4m3fn else (-128.0, 127.0)
    _check(fp8_min == lo and fp8_max == hi, f"{name}: min_8bit / max_8bit must be {lo} / {hi} for {output_q.dtype}")
    masked = masked_m is not None
    _check(output_s.dim() == (3 if masked else 2), f"Expected output_s.dim() == {3 if masked else 2}")
    col_major = output_s.stride(-2) < output_s.stride(-1)
    if not masked and not col_major:       # row-major layouts: identical to the v1 op
        _quant_common(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0, fuse,
                      _SUPPORTED_GROUPS_V2 + (256, 512), name)
        return
    _check(input.numel() > 0, f"{name}: empty input")
    _check(input.is_contiguous(), "input must be contiguous")
    _check(output_q.is_contiguous(), "output_q must be contiguous")
    _check(input.dtype in (torch.bfloat16, torch.float16), f"{name}: input dtype must be Float16 or BFloat16, got {input.dtype}")
    _check(group_size in _SUPPORTED_GROUPS_V2 + (256, 512), f"Unsupported group_size: {group_size}")
    hidden = output_q.size(-1)
    tokens = output_q.size(-2)
    _check(hidden % group_size == 0, "hidden size must be divisible by group_size")
    _check(input.size(-1) == hidden * (2 if fuse else 1) and input.numel() == output_q.numel() * (2 if fuse else 1),
           "input / output_q shapes do not match")
    if masked:
        _check(col_major and bool(scale_ue8m0) and fuse,
               f"{name}: masked_m needs column-major UE8M0 scales and fuse_silu_and_mul (the only masked form of the op)")
        _check(input.dim() == 3 and output_q.dim() == 3, "masked layout: input and output_q must be [experts, tokens, hidden]")
        experts = input.size(0)
        _check(masked_m.is_cuda and masked_m.dtype == torch.int32 and masked_m.is_contiguous() and masked_m.numel() == experts,
               "masked_m must be a contiguous CUDA int32 tensor with one entry per expert")
        _check(output_s.size(0) == experts, "output_s must have one slab per expert")
        expert_stride = output_s.stride(0)
    else:
        experts, expert_stride = 1, 0
        tokens = output_q.numel() // hidden
    if scale_ue8m0:
        _check(output_s.dtype == torch.int32, "UE8M0 column-major scales must be packed int32")
        layout = _lib.SCALE_UE8M0_COL_PACKED
        _check(output_s.size(-1) * 4 >= hidden // group_size, "output_s has too few packed columns")
    else:
        _check(output_s.dtype == torch.float32, "output_s must be float32 unless scale_ue8m0")
        layout = _lib.SCALE_F32_COL_MAJOR
        _check(output_s.size(-1) >= hidden // group_size, "output_s has too few columns")
    _check(output_s.stride(-2) == 1, "column-major output_s must have unit stride along tokens")
    _check(output_s.size(-2) >= tokens, "output_s has fewer rows than tokens")
    lib = _lib.load()
    with torch.cuda.device(input.device):
        rc = lib.sglk_per_token_group_quant_8bit_v2(
            input.data_ptr(), output_q.data_ptr(), output_s.data_ptr(), experts, tokens, hidden, group_size,
            _IN_DTYPES[input.dtype], _OUT_DTYPES[output_q.dtype], eps, fp8_min, fp8_max, int(bool(scale_ue8m0)),
            layout, output_s.stride(-1), expert_stride, int(fuse), masked_m.data_ptr() if masked else None,
            _stream_ptr(input))
    _lib.check(rc, name)


_GEMM_OUT = {torch.bfloat16: _lib.BF16, torch.float16: _lib.F16}


def _gemm_checks(mat_a, mat_b, scales_a, scales_b, out_dtype):
    _check(mat_a.dim() == 2 and mat_b.dim() == 2, "mat_a and mat_b must be 2-D")
    _check(mat_a.size(1) == mat_b.size(0), "mat_a and mat_b shapes cannot be multiplied")
    _check(mat_a.dtype == torch.float8_e4m3fn and mat_b.dtype == torch.float8_e4m3fn,
           "mat_a and mat_b must be Float8_e4m3fn")
    _check(scales_a.dtype == torch.float32 and scales_b.dtype == torch.float32, "scales must be Float32")
    _check(out_dtype in _GEMM_OUT, "out_dtype must be Half or BFloat16")
    M, K = mat_a.shape
    N = mat_b.size(1)
    _check(K == 0 or mat_a.stride(1) == 1 or M == 0, "mat_a must be a row major tensor")
    _check(K == 0 or mat_b.stride(0) == 1 or N == 0, "mat_b must be a column major tensor")
    _check(K % 16 == 0, "mat_a.size(1) must be a multiple of 16 (TMA row pitch)")
    kb, nb = -(K // -128), -(N // -128)
    _check(scales_a.dim() == 2 and tuple(scales_a.shape) == (M, kb), f"scales_a must be [M, ceil(K/128)] = [{M}, {kb}]")
    _check(scales_b.dim() == 2 and tuple(scales_b.shape) == (kb, nb),
           f"scales_b must be [ceil(K/128), ceil(N/128)] = [{kb}, {nb}]")
    return M, N, K


# Stream-K workspaces: one zero-initialised buffer per (device, stream), grown on demand.  The kernel
# leaves the counter region zeroed, so the buffer is reused by consecutive launches on its stream
# (the same per-process cache idea as the reference's `_get_cache_buf`, python/sgl_kernel/gemm.py:80).
# A buffer that has been handed to a launch is NEVER given back to the allocator: a CUDA graph captured earlier still
# holds its address, and if the memory went to another tensor a replay would meet non-zero tile counters (tiles never
# finished, silently wrong output) and would scribble partials over that tensor.  Replaced buffers are parked in
# `_retired` for the life of the process; growth is geometric so that list stays short.
_workspaces = {}
_retired = []
_workspaces_lock = threading.Lock()


def _get_workspace(device: torch.device, stream_ptr: int, nbytes: int) -> Optional[torch.Tensor]:
    """The cached workspace of (device, stream), or None when one would have to be created inside a CUDA-graph
    capture: its zero-fill would only be recorded, not executed, and an eager launch before the first replay would
    then meet garbage tile counters.  The op falls back to whole-tile scheduling for that call (same results).
    The key is the stream the launch goes to: to get stream-K inside a graph, run the warm-up ON THE CAPTURE STREAM
    (torch.cuda.graph captures on a side stream, not on the stream an earlier eager warm-up used) or call
    `reserve_gemm_workspace(shapes, stream=capture_stream)` before capturing."""
    key = (device.index, stream_ptr)
    with _workspaces_lock:
        ws = _workspaces.get(key)
        if ws is None or ws.numel() < nbytes:
            if torch.cuda.is_current_stream_capturing():
                return None
            size = max(nbytes, 1 << 20)
            if ws is not None:
                size = max(size, 2 * ws.numel())
                _retired.append(ws)          # graphs captured so far keep a valid, still-zeroed buffer
            ws = torch.zeros(size, dtype=torch.uint8, device=device)
            _workspaces[key] = ws
        return ws


def _invalidate_workspace(device: torch.device, stream_ptr: int) -> None:
    """After a failed launch the tile counters may be non-zero: stop using the buffer (the next call gets a fresh zeroed
    one) but keep it alive, for the same reason as above."""
    with _workspaces_lock:
        ws = _workspaces.pop((device.index, stream_ptr), None)
        if ws is not None:
            _retired.append(ws)


def reserve_gemm_workspace(shapes, device=None, stream=None) -> int:
    """Create (or grow), outside any capture, the workspace that fp8_blockwise_scaled_mm launches on `stream` of
    `device` will use, large enough for every (M, N, K) in `shapes`; returns its size in bytes.  Call it with the
    stream a CUDA graph is going to be captured on, so the captured launches can use stream-K scheduling, and with
    the largest shapes the process will see, so the buffer never has to be replaced afterwards."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    lib = _lib.load()
    with torch.cuda.device(dev):
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("reserve_gemm_workspace must be called outside CUDA-graph capture")
        sp = (torch.cuda.current_stream(dev) if stream is None else stream).cuda_stream
        need = 1 << 20
        for M, N, K in shapes:
            need = max(need, lib.sglk_gemm_workspace_bytes(int(M), int(N), int(K), -1, 0, 0),
                       lib.sglk_gemm_workspace_bytes(int(M), int(N), int(K), -1, 0, _lib.STREAMK_FORCE))
        ws = _get_workspace(dev, sp, need)
    return ws.numel()


# Scheduling override for tests and tuning tools (thread-local, Python-side only — the C library keeps no state):
# "auto" = the cost model, "force" = stream-K whenever the shape allows it, "never" = whole tiles.
_sched = threading.local()


class stream_k_mode:
    def __init__(self, mode: str):
        assert mode in ("auto", "force", "never")
        self.mode = mode

    def __enter__(self):
        self.prev = getattr(_sched, "mode", "auto")
        _sched.mode = self.mode
        return self

    def __exit__(self, *exc):
        _sched.mode = self.prev
        return False


def _sched_flags(stream_k=True, group_m: int = 0, split: int = 0) -> int:
    mode = getattr(_sched, "mode", "auto")
    if stream_k == "force":
        mode = "force"
    elif stream_k is False or stream_k == "never":
        mode = "never"
    flags = {"auto": _lib.STREAMK_AUTO, "force": _lib.STREAMK_FORCE, "never": _lib.STREAMK_NEVER}[mode]
    return flags | ((int(group_m) & 0xFF) << _lib.GROUP_M_SHIFT) | ((int(split) & 0xFF) << _lib.SPLIT_SHIFT)


def _gemm(mat_a, mat_b, scales_a, scales_b, out_dtype, variant: int = -1, num_sms: int = 0, stream_k=True, group_m: int = 0,
          split: int = 0, extra_flags: int = 0):
    """fp8_blockwise_scaled_mm, include/sgl_kernel_ops.h:581-586 (allocates and returns C).
    stream_k: True (cost model), "force", or False ("never": whole tiles, no workspace)."""
    _check(mat_a.is_cuda and mat_b.is_cuda and scales_a.is_cuda and scales_b.is_cuda,
           "fp8_blockwise_scaled_mm: tensors must be CUDA tensors")
    M, N, K = _gemm_checks(mat_a, mat_b, scales_a, scales_b, out_dtype)
    out = torch.empty((M, N), dtype=out_dtype, device=mat_a.device)
    if M == 0 or N == 0:
        return out
    lda = mat_a.stride(0) if M > 1 else max(K, 16)
    ldb = mat_b.stride(1) if N > 1 else max(K, 16)
    _check(lda % 16 == 0 and ldb % 16 == 0 and mat_a.data_ptr() % 16 == 0 and mat_b.data_ptr() % 16 == 0,
           "mat_a / mat_b rows must be 16-byte aligned")
    lib = _lib.load()
    flags = _sched_flags(stream_k, group_m, split) | int(extra_flags)
    with torch.cuda.device(mat_a.device):
        stream = _stream_ptr(mat_a)
        ws_ptr, ws_bytes = None, 0
        if stream_k is not False:   # False: no workspace at all (whole tiles, scales read in whatever layout they have)
            need = lib.sglk_gemm_workspace_bytes(M, N, K, variant, num_sms, flags)
            ws = _get_workspace(mat_a.device, stream, need) if need > 0 else None
            if ws is not None:
                ws_ptr, ws_bytes = ws.data_ptr(), ws.numel()
        rc = lib.sglk_fp8_blockwise_scaled_mm_ex(
            mat_a.data_ptr(), mat_b.data_ptr(), scales_a.data_ptr(), scales_b.data_ptr(), out.data_ptr(),
            M, N, K, lda, ldb, out.stride(0), scales_a.stride(0), scales_a.stride(1), scales_b.stride(0),
            scales_b.stride(1), _GEMM_OUT[out_dtype], variant, num_sms, flags, ws_ptr, ws_bytes, stream)
        if rc != 0 and ws_ptr is not None:
            _invalidate_workspace(mat_a.device, stream)
    _lib.check(rc, "fp8_blockwise_scaled_mm")
    return out


def _gemm_allgather(mat_a, mat_b, scales_a, scales_b, out_full, col_lo: int, peer_ptrs, variant: int = -1,
                    num_sms: int = 0, multicast_ptr: int = 0):
    """Fused GEMM -> all-gather: this rank's [M, N_local] panel is written at columns [col_lo, col_lo+N_local)
    of `out_full` ([M, N_total], this rank's symmetric buffer) AND at the same place of every peer buffer
    (`peer_ptrs`: base device pointers of the peers' buffers, peer-mapped into this process)."""
    _check(mat_a.is_cuda and mat_b.is_cuda and out_full.is_cuda, "fp8_blockwise_scaled_mm: tensors must be CUDA tensors")
    M, N, K = _gemm_checks(mat_a, mat_b, scales_a, scales_b, out_full.dtype)
    _check(out_full.dim() == 2 and out_full.size(0) == M and out_full.stride(1) == 1, "out_full must be [M, N_total] row-major")
    _check(0 <= col_lo and col_lo + N <= out_full.size(1), "panel does not fit in out_full")
    _check(len(peer_ptrs) <= 7, "at most 8 GPUs")
    if M == 0 or N == 0:
        return out_full
    esz = out_full.element_size()
    ldc = out_full.stride(0)
    lda = mat_a.stride(0) if M > 1 else max(K, 16)
    ldb = mat_b.stride(1) if N > 1 else max(K, 16)
    lib = _lib.load()
    with torch.cuda.device(mat_a.device):
        stream = _stream_ptr(mat_a)
        ws_ptr, ws_bytes = None, 0
        ag_variant = variant if variant >= 0 else (3 if (M > 128 and N >= 256) else 0)   # the entry point's own default
        need = lib.sglk_gemm_workspace_bytes(M, N, K, ag_variant, num_sms, 0)
        ws = _get_workspace(mat_a.device, stream, need) if need > 0 else None
        if ws is not None:
            ws_ptr, ws_bytes = ws.data_ptr(), ws.numel()
        import ctypes

        peers = (ctypes.c_void_p * max(1, len(peer_ptrs)))(*[int(p) + col_lo * esz for p in peer_ptrs])
        rc = lib.sglk_fp8_blockwise_scaled_mm_allgather(
            mat_a.data_ptr(), mat_b.data_ptr(), scales_a.data_ptr(), scales_b.data_ptr(),
            out_full.data_ptr() + col_lo * esz, peers, len(peer_ptrs),
            (int(multicast_ptr) + col_lo * esz) if multicast_ptr else None, M, N, K, lda, ldb, ldc,
            scales_a.stride(0), scales_a.stride(1), scales_b.stride(0), scales_b.stride(1),
            _GEMM_OUT[out_full.dtype], variant, num_sms, ws_ptr, ws_bytes, stream)
    _lib.check(rc, "fp8_blockwise_scaled_mm (fused all-gather)")
    return out_full


ROWS_PER_COUNTER = 256     # one arrival counter per 256-row m-block of the paired prefill tiles


def counter_increments(hidden: int) -> int:
    """How much one call of `_quant_bcast` adds to the counter of every 256-row block it covers."""
    return ROWS_PER_COUNTER * hidden // 16384


def _quant_bcast(x, q_ptr: int, s_ptr: int, counters_ptr: int, multicast: bool, eps: float = 1e-10):
    """Row-sharded group quantisation (include/sglk_fp8.h: sglk_per_token_group_quant_fp8_bcast): quantise `x`
    ([rows, hidden] bf16/f16, this rank's rows) to e4m3 codes + row-major fp32 scales written at the raw addresses
    `q_ptr` / `s_ptr` (NVLS multicast addresses when `multicast`), and bump the arrival counters at `counters_ptr`."""
    _check(x.is_cuda and x.dim() == 2 and x.is_contiguous(), "x must be a contiguous [rows, hidden] CUDA tensor")
    _check(x.dtype in (torch.bfloat16, torch.float16), "x must be bf16 or f16")
    rows, hidden = x.shape
    if rows == 0:
        return
    lib = _lib.load()
    with torch.cuda.device(x.device):
        rc = lib.sglk_per_token_group_quant_fp8_bcast(
            x.data_ptr(), int(q_ptr), int(s_ptr), int(counters_ptr), rows, hidden, ROWS_PER_COUNTER,
            _lib.BF16 if x.dtype == torch.bfloat16 else _lib.F16, float(eps), 1 if multicast else 0, _stream_ptr(x))
    _lib.check(rc, "per_token_group_quant_fp8 (row-sharded broadcast)")


def _gemm_gated(mat_a, mat_b, scales_a, scales_b, out_dtype, counters, target: int, panel_rank: int, panel_world: int,
                num_sms: int = 0, whole_panels: bool = False, wait_polls: int = 0):
    """fp8_blockwise_scaled_mm whose activation rows arrive during the launch (include/sglk_fp8.h:
    sglk_fp8_blockwise_scaled_mm_gated).  `counters`: uint32/int32 tensor, one entry per 256 rows of mat_a.
    wait_polls: 0 = wait for late rows as long as it takes; n > 0 = trap (sticky launch failure, the CUDA context is lost)
    after n polling rounds of ~1-2 us."""
    _check(mat_a.is_cuda and mat_b.is_cuda and scales_a.is_cuda and scales_b.is_cuda and counters.is_cuda,
           "fp8_blockwise_scaled_mm: tensors must be CUDA tensors")
    M, N, K = _gemm_checks(mat_a, mat_b, scales_a, scales_b, out_dtype)
    _check(counters.element_size() == 4 and counters.numel() >= -(M // -ROWS_PER_COUNTER), "one 32-bit counter per 256 rows")
    out = torch.empty((M, N), dtype=out_dtype, device=mat_a.device)
    if M == 0 or N == 0:
        return out
    lda = mat_a.stride(0) if M > 1 else max(K, 16)
    ldb = mat_b.stride(1) if N > 1 else max(K, 16)
    lib = _lib.load()
    with torch.cuda.device(mat_a.device):
        rc = lib.sglk_fp8_blockwise_scaled_mm_gated(
            mat_a.data_ptr(), mat_b.data_ptr(), scales_a.data_ptr(), scales_b.data_ptr(), out.data_ptr(), M, N, K, lda, ldb,
            out.stride(0), scales_a.stride(0), scales_a.stride(1), scales_b.stride(0), scales_b.stride(1),
            _GEMM_OUT[out_dtype], num_sms, counters.data_ptr(), int(target) & 0xFFFFFFFF, int(panel_rank), int(panel_world),
            1 if whole_panels else 0, int(wait_polls), _stream_ptr(mat_a))
    _lib.check(rc, "fp8_blockwise_scaled_mm (gated)")
    return out


def _gemm_quant(mat_a, mat_b, scales_a, scales_b, round_dtype, eps, fuse_silu_and_mul, column_major_scales):
    """fp8_blockwise_scaled_mm followed by sgl_per_token_group_quant_8bit (group 128, e4m3; optionally silu(gate)*up in
    between, gate/up interleaved by 128-column blocks) as ONE kernel: the bf16 intermediate never reaches HBM.
    Returns (codes [M, Nq], scales [M, Nq/128]) — bit-identical to the two (three) separate ops."""
    name = "fp8_blockwise_scaled_mm_quant"
    _check(mat_a.is_cuda and mat_b.is_cuda and scales_a.is_cuda and scales_b.is_cuda, f"{name}: tensors must be CUDA tensors")
    M, N, K = _gemm_checks(mat_a, mat_b, scales_a, scales_b, round_dtype)
    fuse = bool(fuse_silu_and_mul)
    _check(K > 0, f"{name}: K must be positive")
    _check(N % (256 if fuse else 128) == 0, f"{name}: N must be a multiple of {256 if fuse else 128}")
    nq = N // 2 if fuse else N
    groups = nq // 128
    q = torch.empty((M, nq), dtype=torch.float8_e4m3fn, device=mat_a.device)
    if column_major_scales:
        s = torch.empty((groups, M), dtype=torch.float32, device=mat_a.device).t()
    else:
        s = torch.empty((M, groups), dtype=torch.float32, device=mat_a.device)
    if M == 0 or N == 0:
        return q, s
    lda = mat_a.stride(0) if M > 1 else max(K, 16)
    ldb = mat_b.stride(1) if N > 1 else max(K, 16)
    _check(lda % 16 == 0 and ldb % 16 == 0 and mat_a.data_ptr() % 16 == 0 and mat_b.data_ptr() % 16 == 0,
           "mat_a / mat_b rows must be 16-byte aligned")
    lib = _lib.load()
    with torch.cuda.device(mat_a.device):
        stream = _stream_ptr(mat_a)
        ws_ptr, ws_bytes = None, 0
        flags = _sched_flags()
        need = lib.sglk_gemm_workspace_bytes(M, N, K, 13 if M > 128 else 2, 0, flags)
        ws = _get_workspace(mat_a.device, stream, need) if need > 0 else None
        if ws is not None:
            ws_ptr, ws_bytes = ws.data_ptr(), ws.numel()
        rc = lib.sglk_fp8_blockwise_scaled_mm_quant(
            mat_a.data_ptr(), mat_b.data_ptr(), scales_a.data_ptr(), scales_b.data_ptr(), q.data_ptr(), s.data_ptr(),
            M, N, K, lda, ldb, q.stride(0), scales_a.stride(0), scales_a.stride(1), scales_b.stride(0), scales_b.stride(1),
            s.stride(0), s.stride(1), _GEMM_OUT[round_dtype], float(eps), int(fuse), 0, flags, ws_ptr, ws_bytes, stream)
    _lib.check(rc, name)
    return q, s


def _gemm_quant_fake(mat_a, mat_b, scales_a, scales_b, round_dtype, eps, fuse_silu_and_mul, column_major_scales):
    nq = mat_b.size(1) // (2 if fuse_silu_and_mul else 1)
    q = torch.empty((mat_a.size(0), nq), dtype=torch.float8_e4m3fn, device=mat_a.device)
    if column_major_scales:
        return q, torch.empty((nq // 128, mat_a.size(0)), dtype=torch.float32, device=mat_a.device).t()
    return q, torch.empty((mat_a.size(0), nq // 128), dtype=torch.float32, device=mat_a.device)


def _gemm_op(mat_a, mat_b, scales_a, scales_b, out_dtype):
    return _gemm(mat_a, mat_b, scales_a, scales_b, out_dtype)


def _grouped_mm(output, a_ptrs, b_ptrs, out_ptrs, a_scales_ptrs, b_scales_ptrs, a, b, scales_a, scales_b,
                stride_a, stride_b, stride_c, layout_sfa, layout_sfb, problem_sizes, expert_offsets, workspace):
    """fp8_blockwise_scaled_grouped_mm, include/sgl_kernel_ops.h:660-678; semantics tests/test_fp8_blockwise_moe.py:66-171:
    a [sum_m, K] e4m3 (experts stacked), b [E, K, N] e4m3 with K contiguous (the transposed view of [E, N, K]),
    scales_a [sum_m, K/128], scales_b [E, K/128, N/128], problem_sizes int32 [E, 3], expert_offsets int32 [E]."""
    name = "fp8_blockwise_scaled_grouped_mm"
    _check(all(t.is_cuda for t in (output, a, b, scales_a, scales_b, problem_sizes, expert_offsets, workspace)),
           f"{name}: tensors must be CUDA tensors")
    _check(a.dim() == 2 and b.dim() == 3 and output.dim() == 2, "a must be 2-D, b 3-D, output 2-D")
    _check(a.dtype == torch.float8_e4m3fn and b.dtype == torch.float8_e4m3fn, "a and b must be Float8_e4m3fn")
    _check(scales_a.dtype == torch.float32 and scales_b.dtype == torch.float32, "scales must be Float32")
    _check(output.dtype in _GEMM_OUT, "output must be Half or BFloat16")
    _check(problem_sizes.dtype == torch.int32 and expert_offsets.dtype == torch.int32,
           "problem_sizes and expert_offsets must be Int32")
    E, K, N = b.shape
    total_m = a.size(0)
    _check(a.size(1) == K and output.size(0) == total_m and output.size(1) == N, "a / b / output shapes do not match")
    _check(problem_sizes.dim() == 2 and problem_sizes.size(0) == E and problem_sizes.size(1) == 3 and
           problem_sizes.is_contiguous(), "problem_sizes must be a contiguous [num_experts, 3] tensor")
    _check(expert_offsets.numel() >= E and expert_offsets.stride(0) == 1, "expert_offsets must have one entry per expert")
    _check(K == 0 or a.stride(1) == 1 or total_m == 0, "a must be a row major tensor")
    _check(b.stride(1) == 1 and b.stride(0) == N * b.stride(2), "b must be [E, K, N] with K contiguous (column major per expert)")
    _check(output.stride(1) == 1, "output must be row major")
    kb, nb = -(K // -128), -(N // -128)
    _check(scales_a.dim() == 2 and scales_a.size(0) == total_m and scales_a.size(1) == kb, "scales_a must be [sum_m, ceil(K/128)]")
    _check(scales_b.dim() == 3 and tuple(scales_b.shape) == (E, kb, nb), "scales_b must be [E, ceil(K/128), ceil(N/128)]")
    if total_m == 0 or N == 0 or E == 0:
        return
    lib = _lib.load()
    need = lib.sglk_grouped_mm_workspace_bytes(E, total_m)
    ws_ptr = workspace.data_ptr()
    pad = (-ws_ptr) % 16
    _check(workspace.is_contiguous() and workspace.numel() * workspace.element_size() >= need + pad,
           f"{name}: workspace too small ({need + pad} bytes needed)")
    with torch.cuda.device(a.device):
        rc = lib.sglk_fp8_blockwise_scaled_grouped_mm(
            a.data_ptr(), b.data_ptr(), scales_a.data_ptr(), scales_b.data_ptr(), output.data_ptr(),
            expert_offsets.data_ptr(), problem_sizes.data_ptr(), E, total_m, N, K,
            a.stride(0) if total_m > 1 else max(K, 16), b.stride(2) if N > 1 else max(K, 16), output.stride(0),
            scales_a.stride(0), scales_a.stride(1), scales_b.stride(0), scales_b.stride(1), scales_b.stride(2),
            _GEMM_OUT[output.dtype], 0, ws_ptr + pad, workspace.numel() * workspace.element_size() - pad, _stream_ptr(a))
    _lib.check(rc, name)


def _token_quant(input, output_q, output_s):
    """at::native::xpu::sgl_per_token_quant_fp8, src/sycl/per_token_quant_fp8.cpp:201-306 (checks :202-222)."""
    name = "sgl_per_token_quant_fp8"
    _check(input.is_cuda and output_q.is_cuda and output_s.is_cuda, f"{name}: tensors must be CUDA tensors")
    _check(input.is_contiguous() and output_q.is_contiguous() and output_s.is_contiguous(), "tensors must be contiguous")
    _check(input.dim() == 2, "input must be 2D [num_tokens, hidden_dim]")
    _check(input.dtype in (torch.bfloat16, torch.float16), "input must be BFloat16/Half tensor")
    _check(output_q.dtype == torch.float8_e4m3fn, "output_q must be Float8_e4m3fn tensor")
    _check(output_s.dtype == torch.float32, "output_s must be Float tensor")
    tokens, hidden = input.shape
    _check(output_q.shape == input.shape, f"output_q must have same shape as input, got {tuple(output_q.shape)}")
    _check(output_s.numel() == tokens, f"output_s must have numel == num_tokens ({tokens}), got {output_s.numel()}")
    _check(hidden % 4 == 0, f"hidden_dim must be divisible by 4, got {hidden}")
    if tokens == 0 or hidden == 0:
        return
    lib = _lib.load()
    with torch.cuda.device(input.device):
        rc = lib.sglk_per_token_quant_fp8(input.data_ptr(), output_q.data_ptr(), output_s.data_ptr(), tokens, hidden,
                                          _IN_DTYPES[input.dtype], _stream_ptr(input))
    _lib.check(rc, name)


_library.impl("sgl_per_token_quant_fp8", _token_quant, "CUDA")
_library.impl("fp8_blockwise_scaled_grouped_mm", _grouped_mm, "CUDA")
_library.impl("sgl_per_token_group_quant_8bit", _quant_v1, "CUDA")
_library.impl("sgl_per_token_group_quant_8bit_v2", _quant_v2, "CUDA")
_library.impl("fp8_blockwise_scaled_mm", _gemm_op, "CUDA")
_library.impl("fp8_blockwise_scaled_mm_quant", _gemm_quant, "CUDA")
_library.impl("fp8_blockwise_scaled_mm_quant", _gemm_quant_fake, "Meta")


# Fake (meta) kernels so the ops trace under torch.compile / FakeTensor, as the reference README asks of
# every op (README.md:84-85).
def _quant_fake(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0):
    return None


def _quant_v2_fake(input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0,
                   fuse_silu_and_mul, masked_m):
    return None


def _gemm_fake(mat_a, mat_b, scales_a, scales_b, out_dtype):
    return torch.empty((mat_a.size(0), mat_b.size(1)), dtype=out_dtype, device=mat_a.device)


def _grouped_fake(*args):
    return None


_library.impl("sgl_per_token_quant_fp8", lambda input, output_q, output_s: None, "Meta")
_library.impl("fp8_blockwise_scaled_grouped_mm", _grouped_fake, "Meta")
_library.impl("sgl_per_token_group_quant_8bit", _quant_fake, "Meta")
_library.impl("sgl_per_token_group_quant_8bit_v2", _quant_v2_fake, "Meta")
_library.impl("fp8_blockwise_scaled_mm", _gemm_fake, "Meta")
