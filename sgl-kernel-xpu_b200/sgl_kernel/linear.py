"""Host-side callers of the two ops: the allocation helpers every caller of the reference writes around
`sgl_per_token_group_quant_8bit` (tests/test_per_token_group_quant_8bit.py:86-164 — the same helper
lives in SGLang's fp8_kernel.py), and the FP8 block-quantized linear layer that chains quant -> GEMM.
Device memory, streams and allocation are PyTorch's; all arithmetic is in libsglk_sm100a.so.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .gemm import fp8_blockwise_scaled_mm, sgl_per_token_group_quant_8bit

FP8_MAX = 448.0  # torch.finfo(torch.float8_e4m3fn).max


def _round_up(v: int, to: int) -> int:
    return -(-v // to) * to


def _token_major_view(lead, tokens: int, tokens_alloc: int, cols: int, dtype, device) -> torch.Tensor:
    """[*lead, tokens, cols] tensor whose TOKEN index is the unit-stride one ("column-major scales"): storage is
    [*lead, cols, tokens_alloc], seen through swapped last axes and cut back to the tokens that exist."""
    store = torch.empty(tuple(lead) + (cols, tokens_alloc), dtype=dtype, device=device)
    return store.transpose(-1, -2).narrow(-2, 0, tokens)


def create_per_token_group_quant_fp8_output_scale(
    x_shape,
    device,
    group_size: int,
    column_major_scales: bool,
    scale_tma_aligned: bool,
    scale_ue8m0: bool,
) -> torch.Tensor:
    """Allocate the scale tensor of `sgl_per_token_group_quant_8bit` for codes of shape `x_shape`, in one of the four
    layouts callers of the reference use (contract: tests/test_per_token_group_quant_8bit.py:86-126):

      row-major fp32              [..., tokens, groups]                      (neither flag)
      column-major fp32           same shape, token stride 1                 (column_major_scales)
      ... TMA aligned             token extent of the storage padded to 4    (+ scale_tma_aligned)
      packed UE8M0                int32 words of four exponent bytes, [..., tokens, ceil(groups/4)], token stride 1,
                                  storage padded to 4 tokens; group size 128 (+ scale_ue8m0)"""
    *lead, tokens, hidden = tuple(x_shape)
    if scale_ue8m0:
        if not (column_major_scales and scale_tma_aligned):
            raise AssertionError("UE8M0 scales are column-major and TMA aligned")
        words = _round_up(hidden // 128, 4) // 4
        return _token_major_view(lead, tokens, _round_up(tokens, 4), words, torch.int32, device)
    groups = hidden // group_size
    if not column_major_scales:
        return torch.empty((*lead, tokens, groups), dtype=torch.float32, device=device)
    return _token_major_view(lead, tokens, _round_up(tokens, 4) if scale_tma_aligned else tokens, groups, torch.float32, device)


_CODE_RANGE = {torch.float8_e4m3fn: (-FP8_MAX, FP8_MAX), torch.int8: (-128.0, 127.0)}


def per_token_group_quant_fp8(
    x: torch.Tensor,
    group_size: int = 128,
    eps: float = 1e-10,
    column_major_scales: bool = False,
    scale_tma_aligned: bool = False,
    scale_ue8m0: bool = False,
    fuse_silu_and_mul: bool = False,
    masked_m: Optional[torch.Tensor] = None,
    enable_v2: Optional[bool] = None,
    dtype: torch.dtype = torch.float8_e4m3fn,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Allocating form of the op: returns (codes, scales) for `x` — the helper every caller of the reference wraps
    around the in-place op (same keywords as tests/test_per_token_group_quant_8bit.py:129-164).  With
    `fuse_silu_and_mul` the input is [..., 2*hidden] (gate | up) and the codes are [..., hidden]; that form and
    `masked_m` exist only in the v2 op, which is then selected unless the caller says otherwise."""
    if x.shape[-1] % group_size:
        raise AssertionError("the last dimension of `x` must be divisible by `group_size`")
    if not x.is_contiguous():
        raise AssertionError("`x` is not contiguous")
    hidden = x.shape[-1] // 2 if fuse_silu_and_mul else x.shape[-1]
    codes_shape = tuple(x.shape[:-1]) + (hidden,)
    codes = torch.empty(codes_shape, dtype=dtype, device=x.device)
    scales = create_per_token_group_quant_fp8_output_scale(codes_shape, x.device, group_size, column_major_scales,
                                                           scale_tma_aligned, scale_ue8m0)
    if x.shape[0] == 0:
        return codes, scales
    lo, hi = _CODE_RANGE[dtype]
    if enable_v2 is None and fuse_silu_and_mul:
        enable_v2 = True
    sgl_per_token_group_quant_8bit(x, codes, scales, group_size, eps, lo, hi, scale_ue8m0, fuse_silu_and_mul, masked_m,
                                   enable_v2)
    return codes, scales


class Fp8BlockLinear:
    """y = x @ W^T with W stored as e4m3 codes [N, K] and one fp32 scale per 128x128 block
    [ceil(N/128), ceil(K/128)] — the layer SGLang builds on the two ops.  `forward` quantises the bf16
    activations per 1x128 group (row-major scales) and calls the blockwise GEMM.  Weights stay resident in HBM; nothing is re-laid-out per call."""

    def __init__(self, weight_q: torch.Tensor, weight_scale: torch.Tensor, out_dtype: torch.dtype = torch.bfloat16):
        assert weight_q.dtype == torch.float8_e4m3fn and weight_q.dim() == 2 and weight_q.is_contiguous()
        n, k = weight_q.shape
        assert tuple(weight_scale.shape) == (-(n // -128), -(k // -128)), "weight_scale must be [N/128, K/128]"
        self.weight_q = weight_q
        self.weight_scale = weight_scale.contiguous()
        self.out_dtype = out_dtype
        self.out_features, self.in_features = n, k

    def quantize_input(self, x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        # row-major scales (the op's default layout): contiguous along K blocks, which is what the GEMM's
        # promotion warps read fastest (one 16-byte load per four K blocks) and what the quant kernel writes fastest
        return per_token_group_quant_fp8(x, 128, column_major_scales=False, enable_v2=False)

    def forward_quantized(self, x_q: torch.Tensor, x_s: torch.Tensor) -> torch.Tensor:
        # mat_b = [K, N] column-major view of the [N, K] codes; scales_b = [K/128, N/128] view
        return fp8_blockwise_scaled_mm(x_q, self.weight_q.t(), x_s, self.weight_scale.t(), self.out_dtype)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1])
        x_q, x_s = self.quantize_input(x2)
        return self.forward_quantized(x_q, x_s).reshape(*lead, self.out_features)

    __call__ = forward


class HostStreamedFp8Linear:
    """y_host = layer(x_host) for a stream of batches living in pinned HOST memory.

    Three CUDA streams (host->device, compute, device->host) and two device buffers per direction keep the
    PCIe copies of batch i+1 / i-1 under the quant+GEMM of batch i, so sustained throughput is bounded by
    the slower PCIe direction instead of the sum of all three.  This is the host-buffer entry point the
    `e2e` figure of bench.py is measured through."""

    def __init__(self, layer: Fp8BlockLinear, max_tokens: int):
        dev = layer.weight_q.device
        self.layer = layer
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
        self.x_dev = [torch.empty((max_tokens, layer.in_features), dtype=torch.bfloat16, device=dev) for _ in range(2)]
        self.y_dev = [None, None]
        self.ev_in = [torch.cuda.Event() for _ in range(2)]
        self.ev_cmp = [torch.cuda.Event() for _ in range(2)]
        self.ev_out = [torch.cuda.Event() for _ in range(2)]
        self.count = 0

    def submit(self, x_host: torch.Tensor, y_host: torch.Tensor) -> None:
        """Enqueue one batch: x_host [M,K] bf16 pinned -> y_host [M,N] pinned.  Asynchronous."""
        assert x_host.is_pinned() and y_host.is_pinned(), "host buffers must be pinned"
        b = self.count & 1
        m = x_host.shape[0]
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_cmp[b])          # x_dev[b] no longer read by the compute of batch i-2
            self.x_dev[b][:m].copy_(x_host, non_blocking=True)
            self.ev_in[b].record(self.s_in)
        with torch.cuda.stream(self.s_cmp):
            self.s_cmp.wait_event(self.ev_in[b])
            self.s_cmp.wait_event(self.ev_out[b])         # y_dev[b] of batch i-2 has left the device
            y = self.layer(self.x_dev[b][:m])
            self.y_dev[b] = y
            self.ev_cmp[b].record(self.s_cmp)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_cmp[b])
            y.record_stream(self.s_out)
            y_host.copy_(y, non_blocking=True)
            self.ev_out[b].record(self.s_out)
        self.count += 1

    def synchronize(self) -> None:
        self.s_out.synchronize()
        self.s_cmp.synchronize()
        self.s_in.synchronize()
