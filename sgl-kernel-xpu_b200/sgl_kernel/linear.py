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


def _ceil_align(x: int, a: int) -> int:
    return (x + a - 1) // a * a


def create_per_token_group_quant_fp8_output_scale(
    x_shape,
    device,
    group_size: int,
    column_major_scales: bool,
    scale_tma_aligned: bool,
    scale_ue8m0: bool,
) -> torch.Tensor:
    """The four scale layouts of the reference's `create_fp8_output_scale`
    (tests/test_per_token_group_quant_8bit.py:86-126): row-major fp32; column-major fp32;
    column-major fp32 with rows padded to 4 ("TMA aligned"); UE8M0 packed int32, column-major."""
    if scale_ue8m0:
        assert column_major_scales and scale_tma_aligned
        *x_batch, x_q_mn, x_q_k = x_shape
        x_s_mn, x_s_k = x_q_mn, x_q_k // 128
        aligned_mn = _ceil_align(x_s_mn, 4)
        aligned_k = _ceil_align(x_s_k, 4)
        return torch.empty((*x_batch, aligned_k // 4, aligned_mn), device=device, dtype=torch.int).transpose(-1, -2)[
            ..., :x_s_mn, :
        ]
    if column_major_scales:
        if scale_tma_aligned:
            aligned_size = _ceil_align(x_shape[-2], 4)
            return torch.empty(
                tuple(x_shape[:-2]) + (x_shape[-1] // group_size, aligned_size), device=device, dtype=torch.float32
            ).transpose(-1, -2)[: x_shape[-2], :]
        return torch.empty(
            (x_shape[-1] // group_size,) + tuple(x_shape[:-1]), device=device, dtype=torch.float32
        ).permute(-1, -2)
    return torch.empty(tuple(x_shape[:-1]) + (x_shape[-1] // group_size,), device=device, dtype=torch.float32)


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
    """Allocate outputs and quantise — `sglapi_per_token_group_quant_fp8`,
    tests/test_per_token_group_quant_8bit.py:129-164."""
    assert x.shape[-1] % group_size == 0, "the last dimension of `x` cannot be divisible by `group_size`"
    assert x.is_contiguous(), "`x` is not contiguous"
    out_shape = (*x.shape[:-1], x.shape[-1] // (2 if fuse_silu_and_mul else 1))
    x_q = torch.empty(out_shape, device=x.device, dtype=dtype)
    x_s = create_per_token_group_quant_fp8_output_scale(
        out_shape, x.device, group_size, column_major_scales, scale_tma_aligned, scale_ue8m0
    )
    if dtype == torch.int8:
        qmin, qmax = -128.0, 127.0
    else:
        qmax = torch.finfo(dtype).max
        qmin = -qmax
    if x.shape[0] > 0:
        if fuse_silu_and_mul and enable_v2 is None:
            enable_v2 = True
        sgl_per_token_group_quant_8bit(
            x, x_q, x_s, group_size, eps, qmin, qmax, scale_ue8m0, fuse_silu_and_mul, masked_m, enable_v2
        )
    return x_q, x_s


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
