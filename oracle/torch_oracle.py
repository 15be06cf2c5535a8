"""CPU oracle for the FP8 block-quantized linear hot path.  TEST INFRASTRUCTURE ONLY.

This module is a restatement, in plain torch-on-CPU, of the reference's own correctness
definitions for the path.  It is *never* imported by the product (`sgl-kernel-xpu_b200/`);
only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it, and only as the checker / the timed CPU baseline.

Parity status: PINNED.  `tests/golden/make_golden.py` executes the reference's own functions
(extracted by AST from `/root/reference/tests/*.py`, because those modules import
`sgl_kernel`/`torch.xpu` at top level and cannot be imported) and commits their outputs under
`tests/golden/`; `tests/test_oracle.py` checks every function here against those vectors bit
for bit.  The one builder's extension — ragged N in `baseline_scaled_mm` — is flagged below
and is unpinned by the reference (its `group_broadcast` asserts divisibility).

All functions must be evaluated on CPU tensors: torch's CUDA `tensor / python_scalar`
multiplies by the reciprocal, which changes ~58 % of the scales by one ulp.

Every function cites the reference file:line it follows (paths relative to /root/reference).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

FP8_E4M3_MAX = 448.0


def _require_cpu(*ts: torch.Tensor) -> None:
    for t in ts:
        if t.device.type != "cpu":
            raise ValueError("oracle must be evaluated on CPU tensors (see module docstring)")


# ----------------------------------------------------------------------------------------------
# quant half
# ----------------------------------------------------------------------------------------------
def ceil_to_ue8m0(x: torch.Tensor) -> torch.Tensor:
    """tests/test_per_token_group_quant_8bit.py:17-19 — scale -> next power of two."""
    return torch.pow(2.0, torch.ceil(torch.log2(x.abs().clamp(min=1e-10))))


def per_token_group_quant_fp8_ref(
    x: torch.Tensor, group_size: int = 128, eps: float = 1e-10, scale_ue8m0: bool = False
) -> Tuple[torch.Tensor, torch.Tensor]:
    """tests/test_per_token_group_quant_8bit.py:23-40.

    amax over each 1 x group_size group (clamped at eps), scale = amax / 448 (true fp32 division),
    codes = RNE_e4m3(x / scale) with *no* clamp.  Returns (codes [M,K] e4m3fn, scales [M,K/g] f32).
    """
    _require_cpu(x)
    assert x.dim() == 2 and x.size(1) % group_size == 0
    num_tokens, hidden_dim = x.shape
    x_view = x.view(num_tokens, -1, group_size)
    x_amax = x_view.abs().float().amax(dim=2).clamp(min=eps)
    scales = x_amax / FP8_E4M3_MAX
    if scale_ue8m0:
        scales = ceil_to_ue8m0(scales)
    x_quantized = (x_view / scales.unsqueeze(2)).to(torch.float8_e4m3fn)
    return x_quantized.view(num_tokens, hidden_dim), scales


def per_token_group_quant_int8_ref(
    x: torch.Tensor, group_size: int = 128, eps: float = 1e-10
) -> Tuple[torch.Tensor, torch.Tensor]:
    """tests/test_per_token_group_quant_8bit.py:43-54 — int8 twin (clamp +-127, truncating cast)."""
    _require_cpu(x)
    assert x.dim() == 2 and x.size(1) % group_size == 0
    num_tokens, hidden_dim = x.shape
    x_view = x.view(num_tokens, -1, group_size)
    x_amax = x_view.abs().float().amax(dim=2).clamp(min=eps)
    scales = x_amax / 127.0
    x_scaled = (x_view / scales.unsqueeze(2)).clamp(-127.0, 127.0)
    return x_scaled.to(torch.int8).view(num_tokens, hidden_dim), scales


def ue8m0_scale_bytes_ref(x: torch.Tensor, group_size: int, eps: float = 1e-10) -> torch.Tensor:
    """tests/test_per_token_group_quant_8bit.py:399-405 — the reference's only *exact* assertion:
    byte = ceil(log2(max(amax/448, 1e-10))) + 127 per (token, group)."""
    _require_cpu(x)
    m, k = x.shape
    xv = x.view(m, k // group_size, group_size).float()
    amax = xv.abs().amax(dim=2).clamp(min=eps)
    exp_s = torch.ceil(torch.log2((amax / FP8_E4M3_MAX).clamp(min=1e-10)))
    return (exp_s.to(torch.int32) + 127).to(torch.uint8)


def per_token_quant_fp8_ref(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """sgl_per_token_quant_fp8: the kernel's definition (/root/reference/src/sycl/per_token_quant_fp8.cpp:105-123:
    scale = rowmax|x| / 448, inv = scale == 0 ? 0 : 1/scale, q = e4m3(clamp(float(x) * inv, +-448))) written the way
    the reference's torch check does it (tests/test_per_token_quant_fp8.py:17-25: multiply by scale.reciprocal(),
    clamp, .to(float8_e4m3fn)); the zero-row guard is the kernel's (the test never draws a zero row)."""
    _require_cpu(x)
    assert x.dim() == 2
    xf = x.to(torch.float32)
    scale = xf.abs().amax(dim=1) / 448.0
    inv = torch.where(scale == 0, torch.zeros_like(scale), scale.reciprocal()).view(-1, 1)
    q = (xf * inv).clamp(min=-448.0, max=448.0).to(torch.float8_e4m3fn)
    return q, scale


def silu_and_mul_ref(x: torch.Tensor) -> torch.Tensor:
    """tests/test_per_token_group_quant_8bit_v2.py:420-431 (fused path): `silu(primary) * secondary`
    evaluated in the input dtype, i.e. silu rounded to bf16, then the product rounded to bf16,
    before amax/quantisation ("next" row f.1 of SURVEY.md §8)."""
    _require_cpu(x)
    d = x.shape[-1] // 2
    return torch.nn.functional.silu(x[..., :d]) * x[..., d:]


# ----------------------------------------------------------------------------------------------
# weight fabrication (test-side producers)
# ----------------------------------------------------------------------------------------------
def v2_column_major_interleave(s_logical: torch.Tensor) -> torch.Tensor:
    """What the v2 op leaves in a COLUMN-major output_s, as the logical [T, G] tensor a caller indexes.

    The reference kernel writes the scale of (token t, group g) to the logical slot (flat // G, flat % G) with
    flat = t + g * T (/root/reference/src/sycl/per_token_group_quant_8bit_v2.cpp:200-229), i.e. the row-major
    flattening of the logical tensor is the group-major list of scales.  Works on fp32 scales or UE8M0 bytes."""
    T, G = s_logical.shape
    return s_logical.t().contiguous().flatten().reshape(T, G)


def v2_column_major_deinterleave(x_s: torch.Tensor) -> torch.Tensor:
    """The inverse, exactly as the reference's test undoes it before comparing with its row-major oracle
    (/root/reference/tests/test_per_token_group_quant_8bit_v2.py:811-817)."""
    T, G = x_s.shape
    return x_s.contiguous().flatten().reshape(G, T).T.contiguous()


def unpack_ue8m0_words(x_s: torch.Tensor, num_groups: int) -> torch.Tensor:
    """[T, ceil(G/4)] int32 words (any strides) -> [T, G] bytes, little-endian within a word
    (`unpack_match_ref`, tests/test_per_token_group_quant_8bit_v2.py:631-704)."""
    T = x_s.shape[0]
    dense = torch.empty(x_s.numel(), dtype=x_s.dtype).copy_(x_s.reshape(-1))   # size-1 dims keep odd strides through .contiguous()
    b = dense.view(torch.uint8).reshape(T, -1)
    return b[:, :num_groups].contiguous()


def ceil_div(a: int, b: int) -> int:
    return -(a // -b)


def per_block_cast_to_fp8(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """tests/test_es_fp8_blockwise_moe.py:50-62 — 128x128 block quant of a weight matrix."""
    _require_cpu(x)
    assert x.dim() == 2
    m, n = x.shape
    x_padded = torch.zeros((ceil_div(m, 128) * 128, ceil_div(n, 128) * 128), dtype=x.dtype)
    x_padded[:m, :n] = x
    x_view = x_padded.view(-1, 128, x_padded.size(1) // 128, 128)
    x_amax = x_view.abs().float().amax(dim=(1, 3), keepdim=True).clamp(1e-4)
    x_scaled = (x_view * (FP8_E4M3_MAX / x_amax)).to(torch.float8_e4m3fn)
    return (
        x_scaled.view_as(x_padded)[:m, :n].contiguous(),
        (x_amax / FP8_E4M3_MAX).view(x_view.size(0), x_view.size(2)),
    )


# ----------------------------------------------------------------------------------------------
# GEMM half
# ----------------------------------------------------------------------------------------------
def scale_shape(shape, group_shape):
    """tests/test_fp8_blockwise_gemm.py:14-20."""
    assert len(shape) == len(group_shape)
    return tuple(ceil_div(shape[i], group_shape[i]) for i in range(len(group_shape)))


def _group_broadcast(t: torch.Tensor, shape, group_shape) -> torch.Tensor:
    """tests/test_fp8_blockwise_gemm.py:45-56 (`group_broadcast`).  Where the target extent is a
    whole number of 128-groups this is exactly the reference's unsqueeze/expand/flatten; where it is
    not (ragged N, e.g. N=2112 with 17 scale columns) the reference asserts — or, when the extent
    happens to divide, silently infers a group size other than 128 — so the BUILDER'S EXTENSION
    `repeat_interleave(128)[:extent]` is used — unpinned by the reference."""
    for i, s in enumerate(shape):
        if t.shape[i] != s and t.shape[i] != 1:
            if s == t.shape[i] * group_shape[i]:
                t = (
                    t.unsqueeze(i + 1)
                    .expand(*t.shape[: i + 1], s // t.shape[i], *t.shape[i + 1 :])
                    .flatten(i, i + 1)
                )
            else:
                t = t.repeat_interleave(group_shape[i], dim=i).narrow(i, 0, s)
    return t


def baseline_scaled_mm(
    a: torch.Tensor,
    b: torch.Tensor,
    scale_a: torch.Tensor,
    scale_b: torch.Tensor,
    out_dtype: torch.dtype,
    bias: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """tests/test_fp8_blockwise_gemm.py:23-63 — dequantise both operands to fp32 with their
    broadcast block scales, one fp32 `torch.mm`, cast to out_dtype."""
    _require_cpu(a, b, scale_a, scale_b)
    sa = _group_broadcast(scale_a, a.shape, (1, 128))
    sb = _group_broadcast(scale_b, b.shape, (128, 128))
    output = torch.mm(sa * a.to(dtype=torch.float32), sb * b.to(dtype=torch.float32)).to(out_dtype)
    if bias is not None:
        output = output + bias
    return output


def grouped_scaled_mm_ref(a, b, scales_a, scales_b, expert_offsets, problem_sizes, out_dtype):
    """fp8_blockwise_scaled_grouped_mm as its test defines it (tests/test_fp8_blockwise_moe.py:66-171): expert g owns
    rows [off_g, off_g + m_g) of the stacked a / scales_a / output and the g-th [K, N] weight and [K/128, N/128] scales;
    every expert is the plain blockwise GEMM (`baseline_scaled_mm`, :27-51 — same arithmetic as the dense oracle)."""
    _require_cpu(a, b, scales_a, scales_b)
    E = b.shape[0]
    out = torch.zeros(a.shape[0], b.shape[2], dtype=out_dtype)
    for g in range(E):
        off, m = int(expert_offsets[g]), int(problem_sizes[g][0])
        if m == 0:
            continue
        out[off:off + m] = baseline_scaled_mm(a[off:off + m], b[g], scales_a[off:off + m], scales_b[g], out_dtype)
    return out


def to_fp8_round(t: torch.Tensor) -> torch.Tensor:
    """`to_fp8`, tests/test_fp8_blockwise_moe.py:20-24: clamp to the e4m3 range, round to integers, cast."""
    return torch.round(t.clamp(min=-FP8_E4M3_MAX, max=FP8_E4M3_MAX)).to(torch.float8_e4m3fn)


def make_gemm_inputs(M: int, N: int, K: int, seed: int = 0):
    """tests/test_fp8_blockwise_gemm.py:66-80 — the reference's synthetic GEMM inputs:
    codes uniform in +-448, scales randn*1e-3, `mat_b` = [N,K].t(), scales M-major / K-block-major."""
    g = torch.Generator().manual_seed(seed)
    fp8_max = FP8_E4M3_MAX
    a_fp32 = (torch.rand(M, K, dtype=torch.float32, generator=g) - 0.5) * 2 * fp8_max
    a_fp8 = a_fp32.clamp(min=-fp8_max, max=fp8_max).to(torch.float8_e4m3fn)
    b_fp32 = (torch.rand(N, K, dtype=torch.float32, generator=g) - 0.5) * 2 * fp8_max
    b_fp8 = b_fp32.clamp(min=-fp8_max, max=fp8_max).to(torch.float8_e4m3fn).t()
    sa_shape = scale_shape(a_fp8.shape, (1, 128))
    sb_shape = scale_shape(b_fp8.shape, (128, 128))
    scale_a = torch.randn(sa_shape, dtype=torch.float32, generator=g) * 0.001
    scale_b = torch.randn(sb_shape, dtype=torch.float32, generator=g) * 0.001
    scale_a = scale_a.t().contiguous().t()
    scale_b = scale_b.t().contiguous().t()
    return a_fp8, b_fp8, scale_a, scale_b


def make_linear_inputs(M: int, N: int, K: int, seed: int = 42):
    """BASELINE.md §3 config 1 — the realistic chain: randn bf16 activations, randn weights block-cast
    with `per_block_cast_to_fp8`; returns (x_bf16, w_codes.t() [K,N] col-major, w_scales.t() [K/128,N/128])."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g, dtype=torch.float32).to(torch.bfloat16)
    w = torch.randn(N, K, generator=g, dtype=torch.float32)
    wq, ws = per_block_cast_to_fp8(w)
    return x, wq.t(), ws.t()


def fp8_linear_ref(x: torch.Tensor, wq_t: torch.Tensor, ws_t: torch.Tensor, out_dtype=torch.bfloat16):
    """The whole hot path on CPU: oracle quant (group 128) then oracle dequant-matmul."""
    xq, xs = per_token_group_quant_fp8_ref(x, 128)
    return baseline_scaled_mm(xq, wq_t, xs, ws_t, out_dtype)


# ----------------------------------------------------------------------------------------------
# error metrics used by the GEMM parity gate (BASELINE.json: rel-err <= 1e-2, cosine >= 0.9999)
# ----------------------------------------------------------------------------------------------
def rel_err(out: torch.Tensor, ref: torch.Tensor) -> float:
    o, r = out.double().flatten(), ref.double().flatten()
    return float((o - r).norm() / r.norm().clamp(min=1e-30))


def cosine(out: torch.Tensor, ref: torch.Tensor) -> float:
    o, r = out.double().flatten(), ref.double().flatten()
    return float((o @ r) / (o.norm() * r.norm()).clamp(min=1e-30))
