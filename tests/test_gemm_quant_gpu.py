"""GPU parity tests for fp8_blockwise_scaled_mm_quant — the GEMM whose epilogue emits the next layer's e4m3 activations
and 1x128 scales.  Bars:
  * BIT-EXACT against the two (three) separate ops on the same device: fp8_blockwise_scaled_mm -> [silu(gate)*up ->]
    sgl_per_token_group_quant_8bit, which are themselves pinned on the reference's torch code;
  * against the CPU oracle chain (baseline_scaled_mm -> per_token_group_quant_fp8_ref): the dequantised result within
    rel-err 2e-2 / cosine 0.999 of the oracle's own dequantised result (GPU and CPU GEMMs round to bf16 from
    differently ordered fp32 sums, so codes at rounding boundaries may differ by one step)."""
import os

import pytest
import torch

from oracle import torch_oracle as O

import sgl_kernel
from sgl_kernel import _lib, _ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def make_inputs(M, N, K, seed, scale_layout="row"):
    """the reference generator (tests/test_fp8_blockwise_gemm.py:66-80), on CPU, deterministic"""
    a, bt, sa, sb = O.make_gemm_inputs(M, N, K, seed)
    return a, bt, sa, sb


def to_dev(a, b, sa, sb, row_scales):
    a_d = a.to(DEV)
    b_d = b.t().contiguous().to(DEV).t() if b.stride(0) != 1 else b.to(DEV)
    if row_scales:
        sa_d = sa.contiguous().to(DEV)
        sb_d = sb.t().contiguous().to(DEV).t()
    else:
        sa_d = sa.t().contiguous().to(DEV).t()
        sb_d = sb.contiguous().to(DEV)
    return a_d, b_d, sa_d, sb_d


def unfused(a, b, sa, sb, dt, fuse, col_major):
    if fuse:
        # the gate|up mode always runs whole tiles; give its reference the same K-summation order (stream-K would
        # change the fp32 order and with it a few bf16 roundings)
        with _ops.stream_k_mode("never"):
            y = _ops._gemm(a, b, sa, sb, dt, 13 if a.shape[0] > 128 else 2)
    else:
        # the fused op runs the prefill kernel's tiles (13: split-N pairs, 2: single-CTA 256-wide below 129 rows) under
        # the same scheduling decision; its unfused twin is that kernel, not the decode kernel small M defaults to
        y = _ops._gemm(a, b, sa, sb, dt, 13 if a.shape[0] > 128 else 2)
    if fuse:
        M, N = y.shape
        # de-interleave the 128-column blocks back to [gate | up]
        y = y.view(M, N // 256, 2, 128).transpose(1, 2).reshape(M, N).contiguous()
    return sgl_kernel.per_token_group_quant_fp8(y, 128, 1e-10, column_major_scales=col_major, fuse_silu_and_mul=fuse,
                                                enable_v2=True if fuse else False), y


def assert_same(q, s, q_ref, s_ref):
    assert q.shape == q_ref.shape and s.shape == s_ref.shape
    nbad = int((s.contiguous().view(torch.int32) != s_ref.contiguous().view(torch.int32)).sum())
    assert nbad == 0, f"{nbad} / {s.numel()} scales differ"
    nbad = int((q.view(torch.uint8) != q_ref.view(torch.uint8)).sum())
    assert nbad == 0, f"{nbad} / {q.numel()} codes differ"


SHAPES = [(1, 128, 512), (5, 256, 512), (127, 512, 1024), (128, 256, 256), (130, 384, 256), (300, 1024, 512),
          (257, 2304, 1024), (1000, 1280, 7168)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("row_scales,col_major_out", [(True, False), (False, True)])
def test_fused_equals_separate_ops(M, N, K, dt, row_scales, col_major_out):
    a, b, sa, sb = to_dev(*O.make_gemm_inputs(M, N, K, seed=M + N), row_scales)
    q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, dt, column_major_scales=col_major_out)
    (q_ref, s_ref), _ = unfused(a, b, sa, sb, dt, False, col_major_out)
    torch.cuda.synchronize()
    assert (s.stride(0) < s.stride(1)) == col_major_out or s.size(1) == 1 or s.size(0) == 1
    assert_same(q, s, q_ref, s_ref)


def silu_epilogue_available():
    """the silu(gate)*up epilogue is compiled only into the laboratory library (SGLK_LIBRARY=.../libsglk_diag.so): it is
    correct but slower than the separate ops, so the product library does not carry it"""
    a, b, sa, sb = to_dev(*O.make_gemm_inputs(8, 256, 256, seed=1), True)
    try:
        sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.bfloat16, fuse_silu_and_mul=True)
        torch.cuda.synchronize()
        return True
    except RuntimeError:
        return False


def needs_silu_epilogue():
    if not silu_epilogue_available():
        pytest.skip("silu(gate)*up epilogue: laboratory build only")


def test_product_library_rejects_the_silu_epilogue():
    from sgl_kernel import _lib
    if "diag" in os.path.basename(_lib.LIB_PATH):
        pytest.skip("laboratory library loaded")
    a, b, sa, sb = to_dev(*O.make_gemm_inputs(8, 256, 256, seed=1), True)
    with pytest.raises(RuntimeError, match="variant"):
        sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.bfloat16, fuse_silu_and_mul=True)


@pytest.mark.parametrize("M,N,K", [(1, 256, 512), (64, 512, 512), (128, 256, 1024), (130, 768, 256), (300, 1024, 512),
                                   (513, 2304 * 2, 1024), (1000, 2560, 7168)])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_fused_silu_mul_equals_separate_ops(M, N, K, dt):
    needs_silu_epilogue()
    a, b, sa, sb = to_dev(*O.make_gemm_inputs(M, N, K, seed=7 * M + N), True)
    # realistic magnitudes for the activation: outputs of order 1 (the generator's are ~1e-2 .. 1)
    sa = sa * 30.0
    q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, dt, fuse_silu_and_mul=True)
    (q_ref, s_ref), _ = unfused(a, b, sa, sb, dt, True, False)
    torch.cuda.synchronize()
    assert q.shape == (M, N // 2)
    assert_same(q, s, q_ref, s_ref)


def test_interleave_gate_up_round_trip():
    """[gate; up] weights -> interleaved layout -> fused op == silu(x @ gate^T) * (x @ up^T) quantised (separate ops on the
    ORIGINAL stacked layout)."""
    needs_silu_epilogue()
    torch.manual_seed(5)
    M, I, K = 200, 512, 1024
    x = torch.randn(M, K, dtype=torch.bfloat16, device=DEV)
    w = torch.randn(2 * I, K)
    wq, ws = O.per_block_cast_to_fp8(w)
    wq, ws = wq.to(DEV), ws.to(DEV)
    xq, xs = sgl_kernel.per_token_group_quant_fp8(x, 128, enable_v2=False)
    with _ops.stream_k_mode("never"):
        y = _ops._gemm(xq, wq.t(), xs, ws.t(), torch.bfloat16, 13)                            # [M, gate | up]
    q_ref, s_ref = sgl_kernel.per_token_group_quant_fp8(y, 128, fuse_silu_and_mul=True, enable_v2=True)
    wq_i, ws_i = sgl_kernel.interleave_gate_up(wq, ws)
    q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(xq, wq_i.t(), xs, ws_i.t(), torch.bfloat16, fuse_silu_and_mul=True)
    assert_same(q, s, q_ref, s_ref)


@pytest.mark.parametrize("M,N,K", [(5, 256, 512), (130, 384, 256), (300, 1024, 512)])
def test_against_cpu_oracle_chain(M, N, K):
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=3)
    y = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    q_o, s_o = O.per_token_group_quant_fp8_ref(y, 128, 1e-10, False)
    ref = (q_o.float().view(M, N // 128, 128) * s_o.view(M, N // 128, 1)).view(M, N)
    ad, bd, sad, sbd = to_dev(a, b, sa, sb, True)
    q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(ad, bd, sad, sbd, torch.bfloat16)
    out = (q.float().view(M, N // 128, 128) * s.view(M, N // 128, 1)).view(M, N).cpu()
    assert O.rel_err(out, ref) <= 2e-2 and O.cosine(out, ref) >= 0.999
    # scales follow the group amax of the bf16 result: equal unless the GEMMs round the group's largest element apart
    close = torch.isclose(s.cpu(), s_o, rtol=1e-2, atol=0)
    assert float(close.float().mean()) > 0.999
    frac_same = float((q.cpu().view(torch.uint8) == q_o.view(torch.uint8)).float().mean())
    assert frac_same > 0.95, frac_same


def test_stream_k_and_whole_tile_schedules_agree():
    """The fused epilogue runs after the stream-K fix-up, and the plain mode follows the same scheduling decision as
    fp8_blockwise_scaled_mm, so it equals the separate ops under either schedule.  The gate|up mode always runs whole
    tiles (its paired epilogue warps must finish the same tiles), so its reference is the whole-tile GEMM: stream-K
    changes the fp32 summation order and with it a few bf16 roundings."""
    a, b, sa, sb = to_dev(*O.make_gemm_inputs(384, 1280, 7168, seed=11), True)
    with _ops.stream_k_mode("never"):
        q0, s0 = sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.bfloat16)
        (q_ref0, s_ref0), _ = unfused(a, b, sa, sb, torch.bfloat16, False, False)
        (qs_ref, ss_ref), _ = unfused(a, b, sa, sb, torch.bfloat16, True, False)
    with _ops.stream_k_mode("force"):
        q1, s1 = sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.bfloat16)
        (q_ref1, s_ref1), _ = unfused(a, b, sa, sb, torch.bfloat16, False, False)
        silu = silu_epilogue_available()
        if silu:
            qs, ss = sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.bfloat16, fuse_silu_and_mul=True)
        torch.cuda.synchronize()
    assert_same(q0, s0, q_ref0, s_ref0)
    assert_same(q1, s1, q_ref1, s_ref1)
    if silu:
        assert_same(qs, ss, qs_ref, ss_ref)


def test_bench_shape_fused_equals_separate():
    """BASELINE config 5 (M=16384, N=18432, K=7168): whole-tensor bit equality, both modes; the fused op then feeds
    the next GEMM (chain property: down-projection of the fused codes == down-projection of the separate ops' codes)."""
    torch.manual_seed(0)
    M, N, K = 16384, 18432, 7168
    a = ((torch.rand(M, K, device=DEV) - 0.5) * 896).to(torch.float8_e4m3fn)
    bt = ((torch.rand(N, K, device=DEV) - 0.5) * 896).to(torch.float8_e4m3fn)
    sa = torch.rand(M, K // 128, device=DEV) * 1e-3 + 5e-4
    sbt = torch.rand(N // 128, K // 128, device=DEV) * 1e-2 + 5e-3
    for fuse in ((False, True) if silu_epilogue_available() else (False,)):
        q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(a, bt.t(), sa, sbt.t(), torch.bfloat16, fuse_silu_and_mul=fuse)
        (q_ref, s_ref), _ = unfused(a, bt.t(), sa, sbt.t(), torch.bfloat16, fuse, False)
        assert_same(q, s, q_ref, s_ref)
        del q, s, q_ref, s_ref


def test_errors_before_launch():
    a, b, sa, sb = to_dev(*O.make_gemm_inputs(8, 384, 256, seed=1), True)
    with pytest.raises(RuntimeError, match="multiple of 256"):
        sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.bfloat16, fuse_silu_and_mul=True)
    a2, b2, sa2, sb2 = to_dev(*O.make_gemm_inputs(8, 192, 256, seed=1), True)
    with pytest.raises(RuntimeError, match="multiple of 128"):
        sgl_kernel.fp8_blockwise_scaled_mm_quant(a2, b2, sa2, sb2, torch.bfloat16)
    with pytest.raises(RuntimeError, match="Half or BFloat16"):
        sgl_kernel.fp8_blockwise_scaled_mm_quant(a, b, sa, sb, torch.float32)
    q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(a[:0], b, sa[:0], sb, torch.bfloat16)
    assert q.shape == (0, 384) and s.shape == (0, 3)
