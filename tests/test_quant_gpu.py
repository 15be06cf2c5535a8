"""GPU parity tests for sgl_per_token_group_quant_8bit (through torch.ops -> C ABI -> CUDA kernel).
Bar: BIT-EXACT codes and scales against the CPU oracle (BASELINE.json), on the shapes, seeds and
distributions of the reference's tests/test_per_token_group_quant_8bit.py, on the committed golden
vectors, and at BASELINE's full sizes through row-sampled exactness."""
import numpy as np
import pytest
import torch

from conftest import from_bits
from oracle import torch_oracle as O

import sgl_kernel
from sgl_kernel import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
EPS = 1e-10


def run_quant(x_cpu, group, dst=torch.float8_e4m3fn, column_major=False, tma_aligned=False, ue8m0=False, **kw):
    x = x_cpu.to(DEV)
    q, s = sgl_kernel.per_token_group_quant_fp8(x, group, EPS, column_major, tma_aligned, ue8m0, dtype=dst,
                                                enable_v2=kw.pop("enable_v2", False), **kw)
    torch.cuda.synchronize()
    return q.cpu(), s.cpu()


def assert_bits_equal(q, s, q_ref, s_ref):
    assert q.shape == q_ref.shape and q.dtype == q_ref.dtype
    qa, qb = q.view(torch.uint8), q_ref.view(torch.uint8)
    nbad = int((qa != qb).sum())
    assert nbad == 0, f"{nbad} / {qa.numel()} codes differ"
    sa, sb = s.contiguous().view(torch.int32), s_ref.contiguous().view(torch.int32)
    nbad = int((sa != sb).sum())
    assert nbad == 0, f"{nbad} / {sa.numel()} scales differ"


# ---- golden vectors produced by the reference's own python -----------------------------------
GOLDEN_FP8 = [("fp8_bf16_128x1024_g128", 128, torch.bfloat16), ("fp8_bf16_64x512_g64", 64, torch.bfloat16),
              ("fp8_bf16_5x7168_g128", 128, torch.bfloat16), ("fp8_bf16_40x2048_g128", 128, torch.bfloat16),
              ("fp8_bf16_32x256_g32", 32, torch.bfloat16), ("fp8_bf16_3x1024_g256", 256, torch.bfloat16),
              ("fp8_bf16_2x1024_g512", 512, torch.bfloat16), ("fp8_f16_128x1024_g128", 128, torch.float16),
              ("fp8_f32_64x512_g64", 64, torch.float32), ("fp8_bf16_64x512_g64_small", 64, torch.bfloat16),
              ("fp8_bf16_64x512_g64_large", 64, torch.bfloat16), ("fp8_bf16_edge_8x512_g128", 128, torch.bfloat16)]


@pytest.mark.parametrize("name,group,sdt", GOLDEN_FP8)
@pytest.mark.parametrize("column_major,aligned", [(False, False), (True, False), (True, True)])
def test_golden_fp8(quant_golden, name, group, sdt, column_major, aligned):
    x = from_bits(quant_golden[name + ".x"], sdt)
    q, s = run_quant(x, group, column_major=column_major, tma_aligned=aligned)
    assert_bits_equal(q, s, from_bits(quant_golden[name + ".q"], torch.float8_e4m3fn),
                      torch.from_numpy(quant_golden[name + ".s"]))


@pytest.mark.parametrize("name,group,sdt", [("int8_bf16_128x1024_g128", 128, torch.bfloat16),
                                            ("int8_f16_128x1024_g128", 128, torch.float16),
                                            ("int8_bf16_16x4096_g64", 64, torch.bfloat16)])
def test_golden_int8(quant_golden, name, group, sdt):
    x = from_bits(quant_golden[name + ".x"], sdt)
    q, s = run_quant(x, group, dst=torch.int8)
    assert_bits_equal(q, s, torch.from_numpy(quant_golden[name + ".q"].view(np.int8)),
                      torch.from_numpy(quant_golden[name + ".s"]))


@pytest.mark.parametrize("name,group", [("ue8m0_128x1024_g128", 128), ("ue8m0_5x2048_g128", 128), ("ue8m0_32x256_g32", 32)])
def test_golden_ue8m0_row_major_bytes(quant_golden, name, group):
    """tests/test_per_token_group_quant_8bit.py:357-417 — exact byte compare of UE8M0 scales."""
    x = from_bits(quant_golden[name + ".x"], torch.bfloat16).to(DEV)
    m, k = x.shape
    x_s = torch.empty((m, k // group), device=DEV, dtype=torch.uint8)
    x_q = torch.empty((m, k), device=DEV, dtype=torch.float8_e4m3fn)
    torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit.default(x, x_q, x_s, group, EPS, -448.0, 448.0, True)
    assert np.array_equal(x_s.cpu().numpy(), quant_golden[name + ".sbyte"])
    assert np.array_equal(x_q.cpu().view(torch.uint8).numpy(), quant_golden[name + ".q"])


def test_ue8m0_packed_column_major(quant_golden):
    name = "ue8m0_128x1024_g128"
    x = from_bits(quant_golden[name + ".x"], torch.bfloat16)
    q, s = run_quant(x, 128, column_major=True, tma_aligned=True, ue8m0=True)
    assert s.dtype == torch.int32 and s.shape == (128, 2)
    b = s.contiguous().view(torch.uint8).view(128, 2, 4).reshape(128, 8)   # little-endian: byte j = group 4*w+j
    assert np.array_equal(b.numpy(), quant_golden[name + ".sbyte"])
    assert np.array_equal(q.view(torch.uint8).numpy(), quant_golden[name + ".q"])


# ---- the reference's parametrisation (tests/test_per_token_group_quant_8bit.py:278-315), seed 42 ----
REF_CASES = [
    (128, 1024, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (128, 1024, 128, torch.int8, torch.bfloat16, False),
    (64, 512, 64, torch.float8_e4m3fn, torch.bfloat16, False),
    (128, 2048, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (256, 4096, 64, torch.int8, torch.bfloat16, False),
    (128, 1024, 128, torch.float8_e4m3fn, torch.bfloat16, True),
    (128, 1024, 128, torch.float8_e4m3fn, torch.float16, False),
    (128, 1024, 128, torch.int8, torch.float16, False),
    (64, 512, 64, torch.float8_e4m3fn, torch.float16, False),
    (128, 1024, 128, torch.float8_e4m3fn, torch.float32, False),
    (128, 1024, 128, torch.int8, torch.float32, False),
    (64, 512, 64, torch.float8_e4m3fn, torch.float32, False),
    (5120, 7168, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5120, 1536, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5120, 512, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5120, 16384, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5120, 18432, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5120, 2048, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (40960, 2048, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5, 7168, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5, 1536, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5, 16384, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5, 18432, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (5, 2048, 128, torch.float8_e4m3fn, torch.bfloat16, False),
    (40, 2048, 128, torch.float8_e4m3fn, torch.bfloat16, False),
]


def oracle_rows(x_cpu, group, dst, rows=None):
    xs = x_cpu if rows is None else x_cpu[rows].contiguous()
    if dst == torch.int8:
        return O.per_token_group_quant_int8_ref(xs, group, EPS)
    return O.per_token_group_quant_fp8_ref(xs, group, EPS)


@pytest.mark.parametrize("m,k,group,dst,sdt,colmajor", REF_CASES)
def test_reference_parametrisation_bit_exact(m, k, group, dst, sdt, colmajor):
    torch.manual_seed(42)
    x = torch.randn(m, k, dtype=sdt)
    q, s = run_quant(x, group, dst=dst, column_major=colmajor)
    # rows are independent: beyond 8 M elements check an evenly strided 1/8 of the rows exactly
    rows = None if m * k <= (8 << 20) else torch.arange(0, m, 8)
    q_ref, s_ref = oracle_rows(x, group, dst, rows)
    if rows is not None:
        q, s = q[rows], s[rows]
    assert_bits_equal(q, s, q_ref, s_ref)


@pytest.mark.parametrize("m", [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768])
@pytest.mark.parametrize("colmajor", [False, True])
def test_baseline_sweep_k7168(m, colmajor):
    """BASELINE.json config 2: K=7168, group 128, M = 1..32768, row- and column-major scales."""
    torch.manual_seed(0)
    x = torch.randn(m, 7168, dtype=torch.bfloat16)
    q, s = run_quant(x, 128, column_major=colmajor, tma_aligned=colmajor)
    rows = None if m <= 1024 else torch.cat([torch.arange(0, m, max(1, m // 256)), torch.tensor([m - 1])])
    q_ref, s_ref = oracle_rows(x, 128, torch.float8_e4m3fn, rows)
    if rows is not None:
        q, s = q[rows], s[rows]
    assert_bits_equal(q, s, q_ref, s_ref)


@pytest.mark.parametrize("scale", [1e-3, 100.0, 1e-20, 3e4])
def test_magnitude_edge_cases(scale):
    """tests/..8bit.py:339-355 (1e-3, 100) plus sub-eps and near-bf16-max magnitudes."""
    torch.manual_seed(42)
    x = (torch.randn(64, 512) * scale).to(torch.bfloat16)
    q, s = run_quant(x, 64)
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(x, 64, EPS)
    assert_bits_equal(q, s, q_ref, s_ref)
    assert (s > 0).all() and torch.isfinite(s).all()


def test_zero_rows_and_empty():
    x = torch.zeros(3, 256, dtype=torch.bfloat16)
    q, s = run_quant(x, 128)
    assert (q.view(torch.uint8) == 0).all()
    assert torch.equal(s, torch.full((3, 2), EPS) / 448.0)
    q, s = run_quant(torch.zeros(0, 256, dtype=torch.bfloat16), 128)
    assert q.shape == (0, 256) and s.shape == (0, 2)


def test_misaligned_views_use_generic_kernel():
    """An input view that is only 2-byte aligned (and codes 1-byte aligned) still quantises on the
    device, bit-exactly (alignment-agnostic kernel; no CPU fallback)."""
    torch.manual_seed(4)
    x = torch.randn(9, 512, dtype=torch.bfloat16)
    buf = torch.empty(9 * 512 + 8, dtype=torch.bfloat16, device=DEV)
    xv = buf[1:1 + 9 * 512].view(9, 512)
    xv.copy_(x)
    qbuf = torch.empty(9 * 512 + 16, dtype=torch.float8_e4m3fn, device=DEV)
    qv = qbuf[3:3 + 9 * 512].view(9, 512)
    s = torch.empty(9, 4, device=DEV)
    sgl_kernel.sgl_per_token_group_quant_8bit(xv, qv, s, 128, EPS, -448.0, 448.0, enable_v2=False)
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(x, 128, EPS)
    assert_bits_equal(qv.cpu().contiguous(), s.cpu(), q_ref, s_ref)


def test_three_d_input():
    torch.manual_seed(1)
    x = torch.randn(3, 7, 512, dtype=torch.bfloat16)
    xd = x.to(DEV)
    q = torch.empty_like(xd, dtype=torch.float8_e4m3fn)
    s = torch.empty(21, 4, device=DEV)
    sgl_kernel.sgl_per_token_group_quant_8bit(xd, q, s, 128, EPS, -448.0, 448.0, enable_v2=False)
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(x.view(21, 512), 128, EPS)
    assert_bits_equal(q.cpu().view(21, 512), s.cpu(), q_ref, s_ref)


def test_custom_clamp_range():
    """fp8_min/fp8_max are honoured (the reference kernel clamps before converting, :174,188)."""
    torch.manual_seed(2)
    x = torch.randn(16, 256, dtype=torch.bfloat16)
    xd = x.to(DEV)
    q = torch.empty_like(xd, dtype=torch.float8_e4m3fn)
    s = torch.empty(16, 2, device=DEV)
    sgl_kernel.sgl_per_token_group_quant_8bit(xd, q, s, 128, EPS, -224.0, 224.0, enable_v2=False)
    xv = x.view(16, 2, 128)
    scale = xv.abs().float().amax(2).clamp(min=EPS) / 224.0
    ref = (xv / scale.unsqueeze(2)).clamp(-224.0, 224.0).to(torch.float8_e4m3fn).view(16, 256)
    assert_bits_equal(q.cpu(), s.cpu(), ref, scale)


def test_fused_silu_mul_v2():
    """v2 fused path vs its oracle (tests/test_per_token_group_quant_8bit_v2.py:408-444); the
    reference's acceptance rule for codes is `assert_all_close_or_tiny_diff` (:599-628)."""
    torch.manual_seed(11)
    x = torch.randn(64, 2 * 2048, dtype=torch.bfloat16)
    q, s = run_quant(x, 128, fuse_silu_and_mul=True, enable_v2=True)
    y = O.silu_and_mul_ref(x).contiguous()
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(y, 128, EPS)
    torch.testing.assert_close(s, s_ref, rtol=1e-3, atol=1e-7)
    d = (q.view(torch.uint8).to(torch.int16) - q_ref.view(torch.uint8).to(torch.int16)).abs()
    assert int((d >= 2).sum()) == 0 and float((d == 1).sum()) / d.numel() < 0.005


V2_COL_SHAPES = [(1, 128), (4, 384), (16, 1536), (64, 512), (127, 1664), (128, 2048), (512, 7168), (1024, 4096)]


@pytest.mark.parametrize("m,k", V2_COL_SHAPES)
@pytest.mark.parametrize("group", [64, 128])
@pytest.mark.parametrize("aligned", [False, True])
def test_v2_column_major_scales_are_interleaved(m, k, group, aligned):
    """v2 op, column-major fp32 scales: the reference kernel's interleaved addressing
    (per_token_group_quant_8bit_v2.cpp:200-229).  Checked the way the reference's own test does
    (tests/test_per_token_group_quant_8bit_v2.py:807-828: de-interleave, compare with the row-major oracle) — here
    bit for bit — and against the forward restatement in the oracle."""
    torch.manual_seed(m * 7 + k)
    x = torch.randn(m, k, dtype=torch.bfloat16)
    q, s = run_quant(x, group, column_major=True, tma_aligned=aligned, enable_v2=True)
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(x, group, EPS)
    assert_bits_equal(q, O.v2_column_major_deinterleave(s), q_ref, s_ref)
    assert torch.equal(s.contiguous(), O.v2_column_major_interleave(s_ref))


@pytest.mark.parametrize("m,k", V2_COL_SHAPES + [(40, 640), (9, 896)])
@pytest.mark.parametrize("fuse", [False, True])
def test_v2_column_major_packed_ue8m0(m, k, fuse):
    """v2 op, packed UE8M0 column-major scales (+ fused silu*mul): same interleave on the unpacked bytes, unused
    tail bytes of the last word zeroed when G % 4 != 0 (…_v2.cpp:231-248).  The UE8M0 bytes are the one output the
    reference's tests compare exactly."""
    torch.manual_seed(m + k)
    G = k // 128
    x = torch.randn(m, k * (2 if fuse else 1), dtype=torch.bfloat16)
    xd = x.to(DEV)
    q = torch.empty(m, k, dtype=torch.float8_e4m3fn, device=DEV)
    s = sgl_kernel.create_per_token_group_quant_fp8_output_scale((m, k), DEV, 128, True, True, True)
    s.fill_(-1)                                        # so that zeroed tail bytes are seen to be written
    sgl_kernel.sgl_per_token_group_quant_8bit(xd, q, s, 128, EPS, -448.0, 448.0, True, fuse, None, enable_v2=True)
    torch.cuda.synchronize()
    y = O.silu_and_mul_ref(x).contiguous() if fuse else x
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(y, 128, EPS, scale_ue8m0=True)
    bytes_ref = O.ue8m0_scale_bytes_ref(y, 128, EPS)
    got = O.unpack_ue8m0_words(s.cpu(), G)
    assert torch.equal(O.v2_column_major_deinterleave(got), bytes_ref)
    if G % 4:
        words = O.unpack_ue8m0_words(s.cpu(), (G + 3) // 4 * 4)
        # logical row r, last packed word: bytes past G % 4 are zero wherever the row's last group was written
        assert int(words[:, G:(G + 3) // 4 * 4].to(torch.int32).abs().sum()) == 0
    d = (q.cpu().view(torch.uint8).to(torch.int16) - q_ref.view(torch.uint8).to(torch.int16)).abs()
    if fuse:
        assert int((d >= 2).sum()) == 0 and float((d == 1).sum()) / d.numel() < 0.005
    else:
        assert int(d.sum()) == 0


@pytest.mark.parametrize("experts,tokens,hidden,counts", [(4, 16, 2048, [16, 0, 5, 9]), (8, 64, 2048, [64, 1, 0, 33, 64, 7, 12, 63]),
                                                           (2, 128, 1024, [100, 128]), (3, 8, 1152, [8, 3, 0])])
def test_v2_masked_moe_layout(experts, tokens, hidden, counts):
    """v2 op, masked MoE layout (…_v2.cpp:464-563): input [E, T, 2*hidden], only tokens < masked_m[e] are quantised,
    scales go to expert slab e of the 3-D packed UE8M0 column-major tensor with the same interleave over the slab's
    T rows; everything else is left untouched.  (The reference's own masked tests are commented out,
    tests/test_per_token_group_quant_8bit_v2.py:376-399, so this pins the kernel's addressing, restated.)"""
    torch.manual_seed(experts * 100 + tokens)
    G = hidden // 128
    x = torch.randn(experts, tokens, 2 * hidden, dtype=torch.bfloat16)
    masked_m = torch.tensor(counts, dtype=torch.int32)
    q = torch.full((experts, tokens, hidden), 0x55, dtype=torch.uint8, device=DEV).view(torch.float8_e4m3fn)
    s = sgl_kernel.create_per_token_group_quant_fp8_output_scale((experts, tokens, hidden), DEV, 128, True, True, True)
    s.fill_(0x5A5A5A5A)
    sgl_kernel.sgl_per_token_group_quant_8bit(x.to(DEV), q, s, 128, EPS, -448.0, 448.0, True, True, masked_m.to(DEV), enable_v2=True)
    torch.cuda.synchronize()
    qc, sc = q.cpu().view(torch.uint8), s.cpu()
    for e in range(experts):
        n = counts[e]
        y = O.silu_and_mul_ref(x[e]).contiguous()
        q_ref, _ = O.per_token_group_quant_fp8_ref(y, 128, EPS, scale_ue8m0=True)
        bytes_ref = O.ue8m0_scale_bytes_ref(y, 128, EPS)
        d = (qc[e, :n].to(torch.int16) - q_ref[:n].view(torch.uint8).to(torch.int16)).abs()
        assert int((d >= 2).sum()) == 0 and (n == 0 or float((d == 1).sum()) / d.numel() < 0.005)
        assert bool((qc[e, n:] == 0x55).all()), "codes of masked tokens were touched"
        got = O.unpack_ue8m0_words(sc[e], (G + 3) // 4 * 4)            # [T, Gpad] logical bytes of the slab
        expect = torch.full((tokens, (G + 3) // 4 * 4), 0x5A, dtype=torch.uint8)
        for t in range(n):
            for g in range(G):
                flat = t + g * tokens
                expect[flat // G, flat % G] = bytes_ref[t, g]
        # tail-byte zeroing follows the slot the LAST logical group lands in (…_v2.cpp:231-248)
        for t in range(n):
            for g in range(G):
                flat = t + g * tokens
                if flat % G == G - 1 and G % 4:
                    expect[flat // G, G:] = 0
        assert torch.equal(got, expect), f"expert {e}"


# ---- the reference's v2 grid (tests/test_per_token_group_quant_8bit_v2.py:326-402), every axis ----
V2_TOKENS = [1, 4, 16, 64, 127, 128, 512, 1024, 4096, 8192]
V2_HIDDEN = [128, 256, 384, 512, 1024, 1536, 1664, 2048, 4096, 7168, 16384]
V2_LAYOUTS = [dict(column_major=False, tma_aligned=False, ue8m0=False), dict(column_major=True, tma_aligned=False, ue8m0=False),
              dict(column_major=True, tma_aligned=True, ue8m0=False), dict(column_major=True, tma_aligned=True, ue8m0=True)]


def _raw_bytes(t):
    """the elements of t as bytes, in logical order (size-1 dimensions keep arbitrary strides through .contiguous(), which
    view(uint8) then rejects: copy into a fresh dense tensor instead)"""
    return torch.empty(t.shape, dtype=t.dtype).copy_(t).view(torch.uint8)


def reference_v2_input(tokens, hidden):
    """create_per_token_group_quant_test_data (:22-80): seed tokens*10000 + hidden, randn bf16 with a random ~50 % of
    the elements (those where a second randn draw is < 0.001) scaled by 10"""
    g = torch.Generator(device=DEV).manual_seed(tokens * 10000 + hidden)
    x = torch.randn(tokens, hidden, device=DEV, dtype=torch.bfloat16, generator=g)
    x[torch.randn(x.shape, device=DEV, generator=g) < 0.001] *= 10
    return x


@pytest.mark.parametrize("tokens", V2_TOKENS)
@pytest.mark.parametrize("hidden", V2_HIDDEN)
def test_v2_reference_grid(tokens, hidden):
    """Every (tokens, hidden) of the reference's v2 grid, and inside each the group sizes 16/32/64/128 x {e4m3, int8} x
    the four scale layouts (the reference skips UE8M0 with group != 128 and int8 with column-major scales, :760-765):
    1870 op calls.  Bar: codes and scales BIT-EXACT against the CPU oracle — tighter than the reference's own
    assert_all_close_or_tiny_diff / rtol 1e-3 — on every row up to 1 Mi elements and on 128 evenly spaced rows above
    (rows are independent; the interleaved column-major addressing is undone on the whole tensor first)."""
    xd = reference_v2_input(tokens, hidden)
    rows = None if tokens * hidden <= (1 << 20) else torch.linspace(0, tokens - 1, 128).long().unique()
    x_rows = (xd if rows is None else xd[rows.to(DEV)]).cpu().contiguous()
    calls = 0
    for group in (16, 32, 64, 128):
        if hidden % group:
            continue
        G = hidden // group
        oracle = {}
        for dst in (torch.float8_e4m3fn, torch.int8):
            for lay in V2_LAYOUTS:
                if (lay["ue8m0"] and group != 128) or (dst == torch.int8 and lay["column_major"]):
                    continue
                q, s = sgl_kernel.per_token_group_quant_fp8(xd, group, EPS, lay["column_major"], lay["tma_aligned"],
                                                            lay["ue8m0"], dtype=dst, enable_v2=True)
                calls += 1
                key = (dst, lay["ue8m0"])
                if key not in oracle:
                    if dst == torch.int8:
                        oracle[key] = O.per_token_group_quant_int8_ref(x_rows, group, EPS)
                    else:
                        oracle[key] = O.per_token_group_quant_fp8_ref(x_rows, group, EPS, scale_ue8m0=lay["ue8m0"])
                q_ref, s_ref = oracle[key]
                s = s.cpu()
                if lay["ue8m0"]:
                    s = O.v2_column_major_deinterleave(O.unpack_ue8m0_words(s, G))
                    s_ref = O.ue8m0_scale_bytes_ref(x_rows, group, EPS)
                elif lay["column_major"]:
                    s = O.v2_column_major_deinterleave(s)
                q = q.cpu()
                if rows is not None:
                    q, s = q[rows], s[rows]
                what = f"group {group} {dst} {lay}"
                assert torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8)), f"codes differ: {what}"
                assert torch.equal(_raw_bytes(s), _raw_bytes(s_ref)), f"scales differ: {what}"
    assert calls > 0


@pytest.mark.parametrize("tokens", [1, 4, 8, 32, 512, 2048, 6144])
@pytest.mark.parametrize("group_ranks", [8, 16, 32, 48])
def test_v2_reference_grid_fused_silu_mul(tokens, group_ranks):
    """second half of the reference's grid (:362-402): hidden 2048, group 128, packed UE8M0 column-major scales, fused
    silu*mul, no masked layout (num_ranks is unused then, kept as a parameter so the 28 ids match the reference's).
    The fused form is held to the REFERENCE'S tolerant rule (assert_all_close_or_tiny_diff, :599-628: no code off by
    two or more, no sign flips, < 0.5 % off by one — < 4 % on tensors of <= 4096 codes), not to bit-exactness: silu
    is evaluated with the device's exp, the oracle's with the host's.  The UE8M0 scale bytes are compared exactly."""
    hidden = 2048
    g = torch.Generator(device=DEV).manual_seed(tokens * 10000 + hidden)
    xd = torch.randn(tokens, 2 * hidden, device=DEV, dtype=torch.bfloat16, generator=g)
    xd[torch.randn(xd.shape, device=DEV, generator=g) < 0.001] *= 10
    q, s = sgl_kernel.per_token_group_quant_fp8(xd, 128, EPS, True, True, True, fuse_silu_and_mul=True, enable_v2=True)
    y = O.silu_and_mul_ref(xd.cpu()).contiguous()
    q_ref, _ = O.per_token_group_quant_fp8_ref(y, 128, EPS, scale_ue8m0=True)
    bytes_ref = O.ue8m0_scale_bytes_ref(y, 128, EPS)
    got = O.v2_column_major_deinterleave(O.unpack_ue8m0_words(s.cpu(), hidden // 128))
    assert torch.equal(got, bytes_ref), "UE8M0 scale bytes differ"
    d = (q.cpu().view(torch.uint8).to(torch.int16) - q_ref.view(torch.uint8).to(torch.int16)).abs()
    n = d.numel()
    assert int((d >= 2).sum()) == 0          # (a sign flip is a distance of 128)
    tiny = float((d == 1).sum()) / n
    assert tiny < 0.005 or (tiny < 0.04 and n <= 4096)


def test_v2_argument_checks():
    x = torch.randn(8, 256, dtype=torch.bfloat16, device=DEV)
    q = torch.empty(8, 256, dtype=torch.float8_e4m3fn, device=DEV)
    s = torch.empty(8, 2, device=DEV)
    v2 = torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit_v2.default
    with pytest.raises(RuntimeError, match="eps"):
        v2(x, q, s, 128, 1e-6, -448.0, 448.0, False, False, None)
    with pytest.raises(RuntimeError, match="min_8bit"):
        v2(x, q, s, 128, EPS, -224.0, 224.0, False, False, None)
    with pytest.raises(RuntimeError, match="output_s.dim"):
        v2(x, q, s, 128, EPS, -448.0, 448.0, False, False, torch.zeros(1, dtype=torch.int32, device=DEV))
    s3 = torch.empty(1, 8, 2, device=DEV)
    with pytest.raises(RuntimeError, match="masked_m needs"):
        v2(x.view(1, 8, 256), q.view(1, 8, 256), s3, 128, EPS, -448.0, 448.0, False, False,
           torch.zeros(1, dtype=torch.int32, device=DEV))


def test_division_scheme_on_device():
    """Exactness of the 3-op division against __fdiv_rn for every bf16 numerator, on the device."""
    torch.manual_seed(5)
    amax = torch.randn(4096).abs().to(torch.bfloat16).float().clamp(min=EPS)
    amax = torch.cat([amax, torch.tensor([EPS, 1.0, 448.0, 3.0e38, 1.17549435e-38 * 448, 7.0, 0.875])])
    scales = (amax / 448.0).to(DEV)
    for dt in (_lib.BF16, _lib.F16):
        bad = torch.zeros(1, dtype=torch.int64, device=DEV)
        rc = _lib.load().sglk_selftest_division(scales.data_ptr(), scales.numel(), dt, bad.data_ptr(),
                                                   torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        assert int(bad.item()) == 0


def test_error_behaviour():
    x = torch.randn(8, 512, dtype=torch.bfloat16, device=DEV)
    q = torch.empty(8, 512, dtype=torch.float8_e4m3fn, device=DEV)
    s = torch.empty(8, 4, device=DEV)
    f = torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit.default
    with pytest.raises(RuntimeError, match="contiguous"):
        f(x.t(), q, s, 128, EPS, -448.0, 448.0, False)
    with pytest.raises(RuntimeError, match="group_size"):
        f(x, q, s, 100, EPS, -448.0, 448.0, False)
    with pytest.raises(RuntimeError, match="dtype"):
        f(x.to(torch.float64), q, s, 128, EPS, -448.0, 448.0, False)
    with pytest.raises(RuntimeError, match="dtype"):
        f(x, q.view(torch.uint8), s, 128, EPS, -448.0, 448.0, False)
    with pytest.raises(RuntimeError, match="dim"):
        f(x, q, s.view(-1), 128, EPS, -448.0, 448.0, False)
    with pytest.raises(RuntimeError, match="output_s.dim"):     # masked layout wants 3-D scales (…_v2.cpp:739)
        torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit_v2.default(
            x, q, s, 128, EPS, -448.0, 448.0, False, False, torch.zeros(1, device=DEV, dtype=torch.int32))


def test_cuda_graph_capture():
    """No allocation and no sync inside the op: capturable (SURVEY.md §8b threading/streams)."""
    x = torch.randn(64, 1024, dtype=torch.bfloat16, device=DEV)
    q = torch.empty(64, 1024, dtype=torch.float8_e4m3fn, device=DEV)
    s = torch.empty(64, 8, device=DEV)
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, EPS, -448.0, 448.0, enable_v2=False)
    x.copy_(torch.randn(64, 1024, dtype=torch.bfloat16))
    g.replay()
    torch.cuda.synchronize()
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(x.cpu(), 128, EPS)
    assert_bits_equal(q.cpu(), s.cpu(), q_ref, s_ref)


@pytest.mark.parametrize("fuse", [False, True])
def test_jit_entry_points_match_the_ops(fuse):
    """sgl_kernel.jit.* (the reference's JIT twins, jit/__init__.py:34-35) give the AOT ops' bits and the reference's
    return values: (output_q, output_s) from v1, None from v2."""
    from sgl_kernel import jit

    T, H = 48, 512
    x = torch.randn(T, H * (2 if fuse else 1), device="cuda").to(torch.bfloat16)
    mk = lambda: (torch.empty(T, H, dtype=torch.float8_e4m3fn, device="cuda"), torch.empty(T, H // 128, device="cuda"))
    q0, s0 = mk()
    torch.ops.sgl_kernel.sgl_per_token_group_quant_8bit_v2(x, q0, s0, 128, 1e-10, -448.0, 448.0, False, fuse, None)
    q1, s1 = mk()
    assert jit.per_token_group_quant_8bit_v2(x, q1, s1, 128, 1e-10, -448.0, 448.0, False, fuse) is None
    assert torch.equal(q0.view(torch.uint8), q1.view(torch.uint8)) and torch.equal(s0, s1)
    assert jit.per_token_group_quant_8bit_v2(x[:0], q1[:0], s1[:0], 128, 1e-10, -448.0, 448.0) is None
    if not fuse:
        q2, s2 = mk()
        rq, rs = jit.per_token_group_quant_8bit(x, q2, s2, 128, 1e-10, -448.0, 448.0)
        assert rq is q2 and rs is s2
        assert torch.equal(q0.view(torch.uint8), q2.view(torch.uint8)) and torch.equal(s0, s2)
