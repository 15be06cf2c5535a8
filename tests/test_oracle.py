"""CPU suite, part 1: pin the oracle (oracle/torch_oracle.py and oracle/quant_oracle.c) against the
golden vectors produced by the reference's own python (tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from conftest import from_bits
from oracle import c_oracle
from oracle import torch_oracle as O

QUANT_CASES = [
    # name, group, src dtype, kind
    ("fp8_bf16_128x1024_g128", 128, torch.bfloat16, "fp8"),
    ("fp8_bf16_64x512_g64", 64, torch.bfloat16, "fp8"),
    ("fp8_bf16_5x7168_g128", 128, torch.bfloat16, "fp8"),
    ("fp8_bf16_40x2048_g128", 128, torch.bfloat16, "fp8"),
    ("fp8_bf16_32x256_g32", 32, torch.bfloat16, "fp8"),
    ("fp8_bf16_3x1024_g256", 256, torch.bfloat16, "fp8"),
    ("fp8_bf16_2x1024_g512", 512, torch.bfloat16, "fp8"),
    ("fp8_f16_128x1024_g128", 128, torch.float16, "fp8"),
    ("fp8_f32_64x512_g64", 64, torch.float32, "fp8"),
    ("fp8_bf16_64x512_g64_small", 64, torch.bfloat16, "fp8"),
    ("fp8_bf16_64x512_g64_large", 64, torch.bfloat16, "fp8"),
    ("fp8_bf16_edge_8x512_g128", 128, torch.bfloat16, "fp8"),
    ("int8_bf16_128x1024_g128", 128, torch.bfloat16, "int8"),
    ("int8_f16_128x1024_g128", 128, torch.float16, "int8"),
    ("int8_bf16_16x4096_g64", 64, torch.bfloat16, "int8"),
]
_DT = {torch.bfloat16: "bf16", torch.float16: "f16", torch.float32: "f32"}


@pytest.mark.parametrize("name,group,sdt,kind", QUANT_CASES)
def test_torch_oracle_quant_matches_reference(quant_golden, name, group, sdt, kind):
    x = from_bits(quant_golden[name + ".x"], sdt)
    if kind == "fp8":
        q, s = O.per_token_group_quant_fp8_ref(x, group)
        q = q.view(torch.uint8)
    else:
        q, s = O.per_token_group_quant_int8_ref(x, group)
        q = q.view(torch.uint8)
    assert np.array_equal(q.numpy(), quant_golden[name + ".q"].view(np.uint8))
    assert np.array_equal(s.numpy().view(np.uint32), quant_golden[name + ".s"].view(np.uint32))


@pytest.mark.parametrize("name,group,sdt,kind", QUANT_CASES)
def test_c_oracle_quant_matches_reference(quant_golden, name, group, sdt, kind):
    q, s, _ = c_oracle.per_token_group_quant(quant_golden[name + ".x"], _DT[sdt], group, dst_int8=(kind == "int8"))
    assert np.array_equal(q, quant_golden[name + ".q"].view(np.uint8))
    assert np.array_equal(s.view(np.uint32), quant_golden[name + ".s"].view(np.uint32))


@pytest.mark.parametrize("name,group", [("ue8m0_128x1024_g128", 128), ("ue8m0_5x2048_g128", 128), ("ue8m0_32x256_g32", 32)])
def test_ue8m0_matches_reference(quant_golden, name, group):
    x = from_bits(quant_golden[name + ".x"], torch.bfloat16)
    q, s = O.per_token_group_quant_fp8_ref(x, group, scale_ue8m0=True)
    assert np.array_equal(q.view(torch.uint8).numpy(), quant_golden[name + ".q"])
    assert np.array_equal(s.numpy(), quant_golden[name + ".s"])
    assert np.array_equal(O.ue8m0_scale_bytes_ref(x, group).numpy(), quant_golden[name + ".sbyte"])
    # C restatement: exact-integer ceil(log2) agrees with the reference's float formula on these inputs
    qc, sc, bc = c_oracle.per_token_group_quant(quant_golden[name + ".x"], "bf16", group, ue8m0=True)
    assert np.array_equal(qc, quant_golden[name + ".q"])
    assert np.array_equal(sc, quant_golden[name + ".s"])
    assert np.array_equal(bc, quant_golden[name + ".sbyte"])


def test_v2_fused_silu_mul_matches_reference(quant_golden):
    x = from_bits(quant_golden["v2_silu_16x1024_g128.x"], torch.bfloat16)
    y = O.silu_and_mul_ref(x)
    q, s = O.per_token_group_quant_fp8_ref(y.contiguous(), 128)
    assert np.array_equal(q.view(torch.uint8).numpy(), quant_golden["v2_silu_16x1024_g128.q"])
    assert np.array_equal(s.numpy(), quant_golden["v2_silu_16x1024_g128.s"])


GEMM_CASES = ["mm_1x128x512_bf16", "mm_3x256x512_f16", "mm_5x512x1024_bf16", "mm_127x128x512_bf16",
              "mm_128x256x1024_f16", "mm_130x384x256_bf16"]


def _gemm_case(gemm_golden, name):
    a = from_bits(gemm_golden[name + ".a"], torch.float8_e4m3fn)
    bt = from_bits(gemm_golden[name + ".bt"], torch.float8_e4m3fn)
    sa = torch.from_numpy(gemm_golden[name + ".sa"])
    sb = torch.from_numpy(gemm_golden[name + ".sb"])
    odt = torch.bfloat16 if name.endswith("bf16") else torch.float16
    o = from_bits(gemm_golden[name + ".o"], odt)
    return a, bt, sa, sb, odt, o


@pytest.mark.parametrize("name", GEMM_CASES)
def test_torch_oracle_gemm_matches_reference(gemm_golden, name):
    a, bt, sa, sb, odt, o = _gemm_case(gemm_golden, name)
    got = O.baseline_scaled_mm(a, bt.t(), sa, sb, odt)
    # same torch, same fp32 mm: bitwise equal to the reference's own output
    assert torch.equal(got.view(torch.int16), o.view(torch.int16))


@pytest.mark.parametrize("name", GEMM_CASES[:4])
def test_c_oracle_gemm_close_to_reference(gemm_golden, name):
    a, bt, sa, sb, odt, o = _gemm_case(gemm_golden, name)
    got = torch.from_numpy(c_oracle.blockwise_scaled_mm(gemm_golden[name + ".a"], gemm_golden[name + ".bt"],
                                                        gemm_golden[name + ".sa"], gemm_golden[name + ".sb"]))
    # the C version sums in double: equal to the fp32 torch.mm up to the output rounding
    assert O.rel_err(got.to(odt), o) < 4e-3
    assert O.cosine(got, o) > 0.99999


def test_chain_and_weight_cast_match_reference(gemm_golden):
    x = from_bits(gemm_golden["chain_8x256x512.x"], torch.bfloat16)
    w = torch.from_numpy(gemm_golden["chain_8x256x512.w"])
    wq, ws = O.per_block_cast_to_fp8(w)
    assert np.array_equal(wq.view(torch.uint8).numpy(), gemm_golden["chain_8x256x512.wq"])
    assert np.array_equal(ws.numpy(), gemm_golden["chain_8x256x512.ws"])
    out = O.fp8_linear_ref(x, wq.t(), ws.t())
    assert torch.equal(out.view(torch.int16), from_bits(gemm_golden["chain_8x256x512.o"], torch.bfloat16).view(torch.int16))


def test_ragged_n_extension_agrees_on_divisible_prefix():
    """The builder's ragged-N broadcast (N=2112 style) must reduce to the reference's on the columns
    the reference can express."""
    a, b, sa, sb = O.make_gemm_inputs(4, 192, 256, seed=9)       # N=192: 2 scale columns, last covers 64
    full = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    left = O.baseline_scaled_mm(a, b[:, :128], sa, sb[:, :1], torch.bfloat16)
    # (different N can change the BLAS blocking, hence the fp32 summation order: compare to 1 bf16 ulp)
    assert O.rel_err(full[:, :128], left) < 2e-3
    # right 64 columns: pad B to 256 columns with zeros -> divisible case of the reference
    bpad = torch.zeros(256, 256, dtype=torch.float8_e4m3fn)
    bpad[:, :192] = b
    ref = O.baseline_scaled_mm(a, bpad, sa, sb, torch.bfloat16)
    assert O.rel_err(full, ref[:, :192]) < 2e-3 and O.cosine(full, ref[:, :192]) > 0.99999


def test_e4m3_rounding_matches_torch_on_every_bf16():
    bits = (np.arange(1 << 16, dtype=np.uint32) << 16)
    f = bits.view(np.float32)
    t = torch.from_numpy(f.copy()).to(torch.float8_e4m3fn).view(torch.uint8).numpy()
    L = c_oracle.lib()
    mine = np.array([L.oracle_f32_to_e4m3fn(float(v)) for v in f], dtype=np.uint8)
    ok = ~np.isnan(f)
    assert np.array_equal(t[ok], mine[ok])
    # NaN inputs: both give the NaN pattern (sign may differ, see DESIGN.md "unpinned inputs")
    assert np.all((t[~ok] & 0x7F) == 0x7F) and np.all((mine[~ok] & 0x7F) == 0x7F)


def test_division_scheme_exact_on_sampled_scales():
    """The CUDA kernel's 3-op division (x*r, residual, correction) gives the same e4m3 code as RN(x/s)
    for every bf16 numerator; the exhaustive sweep over all bf16 amax (461 M pairs, 0 mismatches) is
    recorded in DESIGN.md — here a slice of it keeps the CPU suite fast."""
    bad = n = 0
    for lo in (0x0000, 0x2E00, 0x3A00, 0x3F80, 0x4100, 0x4700, 0x7F00):
        b, c = c_oracle.check_division_scheme(lo, lo + 0x40)
        bad += b
        n += c
    for lo in (0x0000, 0x0400, 0x2000, 0x3C00, 0x5000, 0x7B00):          # IEEE half numerators
        b, c = c_oracle.check_division_scheme(lo, lo + 0x40, src_f16=True)
        bad += b
        n += c
    for qmax in (127.0, 224.0):                                          # int8 / custom clamp ranges: scale division
        b, c = c_oracle.check_division_scheme(0x3F00, 0x3F40, qmax=qmax)
        bad += b
        n += c
    assert n > 1_000_000 and bad == 0


def test_v2_interleave_matches_the_reference_comment():
    """per_token_group_quant_8bit_v2.cpp:203-212 spells the pattern out for 2 tokens x 3 groups:
    logical [[t0_g0, t1_g0, t0_g1], [t1_g1, t0_g2, t1_g2]]; the test-side inverse is :811-817."""
    s = torch.tensor([[0.0, 1.0, 2.0], [10.0, 11.0, 12.0]])          # s[t, g] = 10 t + g
    inter = O.v2_column_major_interleave(s)
    assert inter.tolist() == [[0.0, 10.0, 1.0], [11.0, 2.0, 12.0]]
    g = torch.Generator().manual_seed(0)
    for T, G in [(1, 1), (5, 3), (4, 56), (127, 13)]:
        r = torch.rand(T, G, generator=g)
        assert torch.equal(O.v2_column_major_deinterleave(O.v2_column_major_interleave(r)), r)
    words = torch.tensor([[0x04030201, 0x00000605]], dtype=torch.int32)
    assert O.unpack_ue8m0_words(words, 6).tolist() == [[1, 2, 3, 4, 5, 6]]


GROUPED = ["grouped0_E4_n256_k512_bf16", "grouped1_E8_n384_k256_f16", "grouped2_E3_n128_k1024_bf16"]


def load_grouped(gg, name):
    odt = torch.bfloat16 if name.endswith("bf16") else torch.float16
    a = from_bits(gg[name + ".a"], torch.float8_e4m3fn)
    bt = from_bits(gg[name + ".bt"], torch.float8_e4m3fn)             # [E, N, K]
    sa = torch.from_numpy(gg[name + ".sa"].copy())
    sbt = torch.from_numpy(gg[name + ".sbt"].copy())                  # [E, N/128, K/128]
    offs = torch.from_numpy(gg[name + ".offsets"].copy())
    sizes = torch.from_numpy(gg[name + ".sizes"].copy())
    o = from_bits(gg[name + ".o"], odt)
    return a, bt, sa, sbt, offs, sizes, o, odt


@pytest.mark.parametrize("name", GROUPED)
def test_grouped_oracle_matches_reference_golden(grouped_golden, name):
    """The grouped restatement against outputs of the reference's own `baseline_scaled_mm` / `to_fp8`
    (tests/test_fp8_blockwise_moe.py), bit for bit."""
    a, bt, sa, sbt, offs, sizes, o, odt = load_grouped(grouped_golden, name)
    got = O.grouped_scaled_mm_ref(a, bt.transpose(1, 2), sa, sbt.transpose(1, 2), offs, sizes, odt)
    assert torch.equal(got.view(torch.int16), o.view(torch.int16))
    x = torch.randn(64, 64, generator=torch.Generator().manual_seed(0)) * 300
    assert torch.equal(O.to_fp8_round(x).view(torch.uint8),
                       torch.round(x.clamp(-448.0, 448.0)).to(torch.float8_e4m3fn).view(torch.uint8))


# ---- sgl_per_token_quant_fp8 (per-row scale) ----------------------------------------------------------------
TOKEN_CASES = [("ptq_128x512_bf16_rand", torch.bfloat16), ("ptq_16x1076_bf16_rand", torch.bfloat16),
               ("ptq_8x1368_f16_rand", torch.float16), ("ptq_4x7168_bf16_randn", torch.bfloat16),
               ("ptq_64x2048_f16_randn", torch.float16), ("ptq_3x20480_bf16_randn", torch.bfloat16)]


@pytest.mark.parametrize("name,sdt", TOKEN_CASES)
def test_token_quant_oracle_matches_reference_golden(token_golden, name, sdt):
    x = from_bits(token_golden[name + ".x"], sdt)
    q, s = O.per_token_quant_fp8_ref(x)
    assert torch.equal(s.view(torch.int32), torch.from_numpy(token_golden[name + ".s"]).view(torch.int32))
    assert torch.equal(q.view(torch.uint8), torch.from_numpy(token_golden[name + ".q"]))


def test_token_quant_oracle_zero_row_follows_the_kernel():
    # the kernel's guard (per_token_quant_fp8.cpp:110): an all-zero row has scale 0 and codes 0, not NaN
    x = torch.zeros(2, 64, dtype=torch.bfloat16)
    x[1, 3] = 1.0
    q, s = O.per_token_quant_fp8_ref(x)
    assert s[0] == 0 and int(q[0].view(torch.uint8).max()) == 0
    assert q[1, 3].float() == 448.0
