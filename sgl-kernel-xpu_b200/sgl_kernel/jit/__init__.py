"""sgl_kernel.jit — the two JIT entry points of the reference that sit on the FP8 block-linear path.

In the reference these compile a SYCL kernel on first use and call it through a thin extern "C" launcher with raw
pointers (python/sgl_kernel/jit/per_token_group_quant_8bit.py:77-200, jit/per_token_group_quant_8bit_v2.py:35-196,
exported at jit/__init__.py:34-35).  This build has nothing to compile at run time — the kernels are already in
libsglk_sm100a.so behind the same kind of C ABI — so the entry points validate like the AOT ops and call that library.
Same names, positional order, defaults and return values as the reference; a CUDA device where it wants an XPU one.
"""
from typing import Optional, Tuple

import torch

from .. import _ops

__all__ = ["per_token_group_quant_8bit", "per_token_group_quant_8bit_v2"]


def per_token_group_quant_8bit(
    input: torch.Tensor,
    output_q: torch.Tensor,
    output_s: torch.Tensor,
    group_size: int,
    eps: float,
    min_8bit: float,
    max_8bit: float,
    scale_ue8m0: bool = False,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Group-quantise `input` into the preallocated `output_q` (int8 / e4m3fn) and `output_s` (scale layout read from
    its dtype and strides); returns the two tensors it filled (reference: jit/per_token_group_quant_8bit.py:153-200)."""
    if input.device.type != "cuda":
        raise RuntimeError("per_token_group_quant_8bit requires a CUDA device")
    _ops._quant_v1(input, output_q, output_s, int(group_size), float(eps), float(min_8bit), float(max_8bit), bool(scale_ue8m0))
    return output_q, output_s


def per_token_group_quant_8bit_v2(
    input: torch.Tensor,
    output_q: torch.Tensor,
    output_s: torch.Tensor,
    group_size: int,
    eps: float,
    min_8bit: float,
    max_8bit: float,
    scale_ue8m0: bool = False,
    fuse_silu_and_mul: bool = False,
    masked_m: Optional[torch.Tensor] = None,
) -> None:
    """The v2 op: optional fused silu(gate)*up, interleaved column-major scales, masked MoE layout (reference:
    jit/per_token_group_quant_8bit_v2.py:118-192).  Returns nothing; an empty input is a no-op as there (:158-160)."""
    assert input.is_contiguous() and output_q.is_contiguous(), "input/output_q must be contiguous"
    assert abs(eps - 1e-10) < 1e-13, "eps must be 1e-10 (matches AOT LOCAL_ABSMAX_ABS)"
    if input.device.type != "cuda":
        raise RuntimeError("per_token_group_quant_8bit_v2 requires a CUDA device")
    if input.numel() // int(group_size) // (2 if fuse_silu_and_mul else 1) == 0:
        return
    _ops._quant_v2(input, output_q, output_s, int(group_size), float(eps), float(min_8bit), float(max_8bit),
                   bool(scale_ue8m0), bool(fuse_silu_and_mul), masked_m)
