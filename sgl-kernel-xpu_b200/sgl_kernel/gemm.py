"""Python surface of the FP8 block-linear path — same names, positional order and defaults as the
reference's python/sgl_kernel/gemm.py:24-31 (`fp8_blockwise_scaled_mm`) and :85-126
(`sgl_per_token_group_quant_8bit` and its legacy aliases)."""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import _ops  # noqa: F401  (registers torch.ops.sgl_kernel.*)


def _get_bool_env_var(name: str, default: str = "false") -> bool:
    # The reference imports sglang.srt.utils.get_bool_env_var lazily (gemm.py:99); sglang is not a
    # dependency here, so the same parsing ("true"/"1", case-insensitive) is done locally.
    return os.getenv(name, default).lower() in ("true", "1")


def fp8_blockwise_scaled_mm(mat_a, mat_b, scales_a, scales_b, out_dtype):
    return torch.ops.sgl_kernel.fp8_blockwise_scaled_mm.default(mat_a, mat_b, scales_a, scales_b, out_dtype)


def sgl_per_token_group_quant_8bit(
    input: torch.Tensor,
    output_q: torch.Tensor,
    output_s: torch.Tensor,
    group_size: int,
    eps: float,
    fp8_min: float,
    fp8_max: float,
    scale_ue8m0: bool = False,
    fuse_silu_and_mul: bool = False,
    masked_m: Optional[torch.Tensor] = None,
    enable_v2: Optional[bool] = None,
) -> None:
    if enable_v2 is None:
        enable_v2 = _get_bool_env_var("SGLANG_PER_TOKEN_GROUP_QUANT_8BIT_V2")

    if enable_v2:
        return torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit_v2.default(
            input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0, fuse_silu_and_mul, masked_m
        )

    assert not fuse_silu_and_mul, "only v2 support fuse_silu_and_mul"
    assert masked_m is None, "only v2 support masked_m"
    torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit.default(
        input, output_q, output_s, group_size, eps, fp8_min, fp8_max, scale_ue8m0
    )


# For legacy usage
sgl_per_token_group_quant_fp8 = sgl_per_token_group_quant_8bit
sgl_per_token_group_quant_int8 = sgl_per_token_group_quant_8bit


def sgl_per_token_quant_fp8(
    input: torch.Tensor,
    output_q: torch.Tensor,
    output_s: torch.Tensor,
) -> None:
    # python/sgl_kernel/gemm.py:236-241
    torch.ops.sgl_kernel.sgl_per_token_quant_fp8.default(input, output_q, output_s)


def fp8_blockwise_scaled_mm_quant(
    mat_a: torch.Tensor,
    mat_b: torch.Tensor,
    scales_a: torch.Tensor,
    scales_b: torch.Tensor,
    round_dtype: torch.dtype = torch.bfloat16,
    eps: float = 1e-10,
    fuse_silu_and_mul: bool = False,
    column_major_scales: bool = False,
):
    """B200 extension: fp8_blockwise_scaled_mm whose epilogue emits the NEXT layer's e4m3 activations and 1x128 scales
    (optionally after silu(gate) * up, gate/up interleaved by 128-column blocks — see `interleave_gate_up`).  Bit-identical
    to fp8_blockwise_scaled_mm -> [silu_and_mul ->] sgl_per_token_group_quant_8bit, without the bf16 round trip."""
    return torch.ops.sgl_kernel.fp8_blockwise_scaled_mm_quant.default(
        mat_a, mat_b, scales_a, scales_b, round_dtype, eps, fuse_silu_and_mul, column_major_scales
    )


def interleave_gate_up(w_q: torch.Tensor, w_s: torch.Tensor):
    """[gate; up] stacked weight ([2I, K] codes, [2I/128, K/128] scales) -> 128-row blocks interleaved
    (gate block 0, up block 0, gate block 1, ...), the layout `fuse_silu_and_mul=True` expects."""
    two_i, k = w_q.shape
    assert two_i % 256 == 0 and w_s.shape[0] == two_i // 128
    nb = two_i // 256
    wq = w_q.view(2, nb, 128, k).transpose(0, 1).reshape(two_i, k).contiguous()
    ws = w_s.view(2, nb, w_s.shape[1]).transpose(0, 1).reshape(two_i // 128, w_s.shape[1]).contiguous()
    return wq, ws
