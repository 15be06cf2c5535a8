"""sgl_kernel — B200 (sm_100a) build of the FP8 block-quantized linear hot path.

Drop-in for the names the reference exports for this path
(python/sgl_kernel/__init__.py:67,75-78); everything else in the reference is out of scope here.
Importing the package registers torch.ops.sgl_kernel.{sgl_per_token_group_quant_8bit,
sgl_per_token_group_quant_8bit_v2, fp8_blockwise_scaled_mm, fp8_blockwise_scaled_grouped_mm,
sgl_per_token_quant_fp8} for dispatch key CUDA.
"""
from .gemm import (
    fp8_blockwise_scaled_mm,
    fp8_blockwise_scaled_mm_quant,
    interleave_gate_up,
    sgl_per_token_group_quant_8bit,
    sgl_per_token_group_quant_fp8,
    sgl_per_token_group_quant_int8,
    sgl_per_token_quant_fp8,
)
from .linear import (
    Fp8BlockLinear,
    HostStreamedFp8Linear,
    create_per_token_group_quant_fp8_output_scale,
    per_token_group_quant_fp8,
)
from ._ops import reserve_gemm_workspace
from .moe import fp8_blockwise_scaled_grouped_mm
from .sharded import ColumnShardedFp8Linear, HostStreamedShardedFp8Linear

__version__ = "0.11.0+b200"

__all__ = [
    "fp8_blockwise_scaled_mm",
    "fp8_blockwise_scaled_mm_quant",
    "interleave_gate_up",
    "fp8_blockwise_scaled_grouped_mm",
    "sgl_per_token_group_quant_8bit",
    "sgl_per_token_group_quant_fp8",
    "sgl_per_token_group_quant_int8",
    "sgl_per_token_quant_fp8",
    "per_token_group_quant_fp8",
    "create_per_token_group_quant_fp8_output_scale",
    "Fp8BlockLinear",
    "HostStreamedFp8Linear",
    "ColumnShardedFp8Linear",
    "HostStreamedShardedFp8Linear",
    "reserve_gemm_workspace",
]
