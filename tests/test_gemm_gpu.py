"""GPU parity tests for fp8_blockwise_scaled_mm (torch.ops -> C ABI -> tcgen05 kernel).
Bar (BASELINE.json): relative error <= 1e-2 and cosine >= 0.9999 against the fp32 dequantised matmul
of the CPU oracle, for bf16 and fp16 outputs; plus the reference test's own
assert_close(rtol=0.02, atol=1) (tests/test_fp8_blockwise_gemm.py:83-85)."""
import pytest
import torch

from conftest import from_bits
from oracle import torch_oracle as O

import sgl_kernel
from sgl_kernel import _lib, _ops

pytestmark = pytest.mark.gpu
DEV = "cuda"
REL_TOL, COS_TOL = 1e-2, 0.9999


def product_variants():
    """Every id the product library can launch (laboratory ids exist only in libsglk_diag.so)."""
    lib = _lib.load()
    return [v for v in range(lib.sglk_gemm_num_variants()) if lib.sglk_gemm_variant_is_product(v)]


PRODUCT = [0, 1, 2, 3, 4, 5, 6, 13, 14, 23, 29, 30, 31, 32]
DECODE = [30, 31, 32]                  # swap-AB decode kernels: single CTA / CTA pair tiles (M <= 256), row-balanced (M <= 128)
DECODE_MAX_M = {30: 256, 31: 256, 32: 128}


def skip_if_too_many_tokens(variant, M):
    if variant in DECODE and M > DECODE_MAX_M[variant]:
        pytest.skip(f"decode kernel {variant}: M <= {DECODE_MAX_M[variant]}")


def test_product_variant_list_is_current():
    assert product_variants() == PRODUCT
    lib = _lib.load()
    lab = [v for v in range(lib.sglk_gemm_num_variants()) if not lib.sglk_gemm_variant_is_product(v)]
    a, b, sa, sb = O.make_gemm_inputs(8, 256, 512, seed=0)
    for v in lab:          # wrong-result diagnostics are unreachable through the product ABI
        with pytest.raises(RuntimeError, match="variant"):
            _ops._gemm(a.to(DEV), b.t().contiguous().to(DEV).t(), sa.to(DEV), sb.to(DEV), torch.bfloat16, v)


def gpu_mm(a, b, sa, sb, odt, variant=-1):
    out = _ops._gemm(a.to(DEV), b.t().contiguous().to(DEV).t(), sa.t().contiguous().to(DEV).t() if sa.shape[0] > 0 else sa.to(DEV),
                     sb.t().contiguous().to(DEV).t(), odt, variant)
    torch.cuda.synchronize()
    return out.cpu()


def check(out, ref, ref_assert_close=True):
    assert out.shape == ref.shape and out.dtype == ref.dtype
    assert torch.isfinite(out.float()).all()
    r, c = O.rel_err(out, ref), O.cosine(out, ref)
    assert r <= REL_TOL and c >= COS_TOL, f"rel_err={r:.3e} cosine={c:.6f}"
    if ref_assert_close:
        torch.testing.assert_close(out, ref, rtol=0.02, atol=1)


GOLDEN = ["mm_1x128x512_bf16", "mm_3x256x512_f16", "mm_5x512x1024_bf16", "mm_127x128x512_bf16",
          "mm_128x256x1024_f16", "mm_130x384x256_bf16"]


@pytest.mark.parametrize("name", GOLDEN)
def test_golden(gemm_golden, name):
    a = from_bits(gemm_golden[name + ".a"], torch.float8_e4m3fn)
    bt = from_bits(gemm_golden[name + ".bt"], torch.float8_e4m3fn)
    sa = torch.from_numpy(gemm_golden[name + ".sa"])
    sb = torch.from_numpy(gemm_golden[name + ".sb"])
    odt = torch.bfloat16 if name.endswith("bf16") else torch.float16
    out = gpu_mm(a, bt.t(), sa, sb, odt)
    check(out, from_bits(gemm_golden[name + ".o"], odt))


SHAPES = [(1, 128, 512), (5, 512, 1024), (127, 256, 512), (128, 128, 7168), (129, 384, 1024), (300, 640, 2048),
          (512, 2112, 1024), (1024, 1024, 512), (64, 4096, 128)]


@pytest.mark.parametrize("variant", PRODUCT)
@pytest.mark.parametrize("M,N,K", SHAPES)
def test_every_variant(variant, M, N, K):
    skip_if_too_many_tokens(variant, M)
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N + K)
    odt = torch.bfloat16 if (M + variant) % 2 == 0 else torch.float16
    ref = O.baseline_scaled_mm(a, b, sa, sb, odt)
    check(gpu_mm(a, b, sa, sb, odt, variant), ref)


# a diagonal slice of the reference grid (tests/test_fp8_blockwise_gemm.py:88-91): every M, N and K
# value of the grid appears; the full 576-case product is bounded by CPU-oracle time, not by the kernel
REF_GRID = [(1, 128, 512), (3, 512, 1024), (5, 1024, 4096), (127, 4096, 8192), (128, 8192, 14080), (512, 14080, 16384),
            (1024, 128, 16384), (4096, 512, 512), (1, 14080, 1024), (3, 8192, 512), (5, 4096, 16384), (127, 1024, 14080),
            (128, 512, 8192), (512, 128, 4096), (1024, 14080, 512), (4096, 8192, 1024), (4096, 128, 8192), (1024, 4096, 4096)]


@pytest.mark.parametrize("M,N,K", REF_GRID)
@pytest.mark.parametrize("odt", [torch.bfloat16, torch.float16])
def test_reference_grid_slice(M, N, K, odt):
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=0)
    rows = None if M * N * K <= (1 << 33) else torch.arange(0, M, M // 128)
    out = gpu_mm(a, b, sa, sb, odt)
    if rows is not None:
        ref = O.baseline_scaled_mm(a[rows], b, sa[rows], sb, odt)
        out = out[rows]
    else:
        ref = O.baseline_scaled_mm(a, b, sa, sb, odt)
    check(out, ref)


def device_baseline_scaled_mm(a, b, sa, sb, odt):
    """the reference test's baseline evaluated where the reference evaluates it — on the device, in fp32
    (tests/test_fp8_blockwise_gemm.py:23-63: broadcast both scale tensors over their 1x128 / 128x128 groups, dequantise,
    one fp32 torch.mm, cast); restated with repeat_interleave"""
    K, N = b.shape
    sa_full = sa.repeat_interleave(128, dim=1)[:, :K]
    sb_full = sb.repeat_interleave(128, dim=0)[:K].repeat_interleave(128, dim=1)[:, :N]
    return torch.mm(sa_full * a.to(torch.float32), sb_full * b.to(torch.float32)).to(odt)


@pytest.mark.parametrize("M", [1, 3, 5, 127, 128, 512, 1024, 4096])
@pytest.mark.parametrize("N", [128, 512, 1024, 4096, 8192, 14080])
@pytest.mark.parametrize("K", [512, 1024, 4096, 8192, 14080, 16384])
@pytest.mark.parametrize("odt", [torch.bfloat16, torch.float16])
def test_reference_grid_full(M, N, K, odt):
    """ALL 576 cases of the reference's own accuracy grid (tests/test_fp8_blockwise_gemm.py:66-91): its generator
    ((rand - 0.5) * 2 * 448 codes, randn * 1e-3 scales made column-major), its baseline on the device, its
    assert_close(rtol=0.02, atol=1) — plus this repo's bars (rel-err <= 1e-2, cosine >= 0.9999)."""
    assert not torch.backends.cuda.matmul.allow_tf32
    g = torch.Generator(device=DEV).manual_seed(M * 1000003 + N * 1009 + K)
    a = ((torch.rand(M, K, device=DEV, generator=g) - 0.5) * 2 * 448.0).clamp(-448.0, 448.0).to(torch.float8_e4m3fn)
    b = ((torch.rand(N, K, device=DEV, generator=g) - 0.5) * 2 * 448.0).clamp(-448.0, 448.0).to(torch.float8_e4m3fn).t()
    sa = (torch.randn(M, K // 128, device=DEV, generator=g) * 0.001).t().contiguous().t()
    sb = (torch.randn(K // 128, N // 128, device=DEV, generator=g) * 0.001).t().contiguous().t()
    ref = device_baseline_scaled_mm(a, b, sa, sb, odt)
    out = sgl_kernel.fp8_blockwise_scaled_mm(a, b, sa, sb, odt)
    torch.testing.assert_close(ref, out, rtol=0.02, atol=1)
    of, rf = out.float(), ref.float()
    rel = float((of - rf).norm() / rf.norm())
    cos = float(torch.nn.functional.cosine_similarity(of.flatten(), rf.flatten(), dim=0))
    assert rel <= REL_TOL and cos >= COS_TOL, f"rel_err={rel:.3e} cosine={cos:.6f}"


def test_quant_then_gemm_chain_config1():
    """BASELINE.json config 1: M=128 N=4096 K=7168 group 128, the realistic chain, GPU vs CPU oracle.
    Quant is bit-exact, so both GEMMs see identical codes and scales."""
    x, wq_t, ws_t = O.make_linear_inputs(128, 4096, 7168, seed=42)
    ref = O.fp8_linear_ref(x, wq_t, ws_t)
    layer = sgl_kernel.Fp8BlockLinear(wq_t.t().contiguous().to(DEV), ws_t.t().contiguous().to(DEV))
    out = layer(x.to(DEV))
    torch.cuda.synchronize()
    check(out.cpu(), ref, ref_assert_close=False)
    torch.testing.assert_close(out.cpu().float(), ref.float(), rtol=0.02, atol=0.05)


def test_chain_golden(gemm_golden):
    x = from_bits(gemm_golden["chain_8x256x512.x"], torch.bfloat16)
    wq = from_bits(gemm_golden["chain_8x256x512.wq"], torch.float8_e4m3fn)
    ws = torch.from_numpy(gemm_golden["chain_8x256x512.ws"])
    layer = sgl_kernel.Fp8BlockLinear(wq.to(DEV), ws.to(DEV))
    out = layer(x.to(DEV)).cpu()
    check(out, from_bits(gemm_golden["chain_8x256x512.o"], torch.bfloat16), ref_assert_close=False)


def test_row_major_activation_scales():
    """scales_a straight from the quant op's default (row-major) layout is accepted too."""
    a, b, sa, sb = O.make_gemm_inputs(200, 512, 1024, seed=3)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    out = _ops._gemm(a.to(DEV), b.t().contiguous().to(DEV).t(), sa.contiguous().to(DEV),
                     sb.contiguous().to(DEV), torch.bfloat16)
    check(out.cpu(), ref)


@pytest.mark.parametrize("M,N,K", [(1, 2112, 7168), (64, 24576, 2048), (200, 512, 1024), (300, 1280, 4096), (1024, 1024, 512),
                                   (129, 384, 1536), (16, 4096, 7168)])
@pytest.mark.parametrize("variant", [-1, 0, 2, 3, 13, 23, 30, 31, 32])
def test_k_contiguous_scales_vector_path(M, N, K, variant, force_stream_k):
    """scales_a row-major (the quant op's default output) and scales_b contiguous along K blocks (the view of a
    [N/128, K/128] weight-scale tensor): the promotion warps' 16-byte-per-4-blocks path, incl. stream-K segments that
    start in the middle of a group of four (K=1536 has 12 blocks: vector path; the others too)."""
    skip_if_too_many_tokens(variant, M)
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + K)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    sad = sa.contiguous().to(DEV)                           # [M, G] row-major
    sbd = sb.t().contiguous().to(DEV).t()                   # (kb, nb) at nb*G + kb
    assert sad.stride(1) == 1 and (sbd.stride(0) == 1 or sbd.shape[0] == 1)
    out = _ops._gemm(a.to(DEV), b.t().contiguous().to(DEV).t(), sad, sbd, torch.bfloat16, variant)
    torch.cuda.synchronize()
    check(out.cpu(), ref)


@pytest.mark.parametrize("M,N,K", [(512, 512, 512), (2048, 4608, 7168), (4099, 2112, 1536), (4099, 7168, 5632), (8192, 2112, 9216)])
@pytest.mark.parametrize("variant", [-1, 0, 3])
def test_scale_relayout_path(M, N, K, variant):
    """M-major scales_a / [G, NB] scales_b (the reference test's layouts, tests/test_fp8_blockwise_gemm.py:23-41) at
    prefill M: with a workspace the op re-lays the scales out K-block-contiguous first.  Same bits as the call that
    is handed K-contiguous scales; the call without a workspace (cp.async-staged scale path, no stream-K) agrees with
    the oracle too.  The first three shapes are below the size where the re-layout pays and do not take it."""
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sa_mmaj = sa.t().contiguous().to(DEV).t()               # (m, kb) at kb*M + m
    sb_rows = sb.contiguous().to(DEV)                       # (kb, nb) at kb*NB + nb
    assert sa_mmaj.stride(0) == 1 and sb_rows.stride(1) == 1
    out = _ops._gemm(ad, bd, sa_mmaj, sb_rows, torch.bfloat16, variant)
    out_nows = _ops._gemm(ad, bd, sa_mmaj, sb_rows, torch.bfloat16, variant, stream_k=False)
    out_vec = _ops._gemm(ad, bd, sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t(), torch.bfloat16, variant)
    torch.cuda.synchronize()
    check(out.cpu(), ref)
    assert torch.equal(out, out_vec)
    check(out_nows.cpu(), ref)


BIG = [("llama3_70b_gate_up", 8192, 28672, 8192), ("dsv3_mlp", 16384, 18432, 7168), ("dsv3_decode_wide", 256, 24576, 7168),
       ("dsv3_decode_ragged", 256, 2112, 7168), ("dsv3_decode_m1", 1, 24576, 7168)]


@pytest.mark.parametrize("name,M,N,K", BIG)
def test_baseline_full_sizes(name, M, N, K):
    """BASELINE.json configs 3-5 at full size: 96 strided rows against the CPU oracle, plus
    size-independent properties — determinism and exact linearity in a power-of-two scale."""
    g = torch.Generator().manual_seed(1)
    a = (torch.rand(M, K, generator=g) - 0.5).mul_(2 * 448).to(torch.float8_e4m3fn)
    bt = (torch.rand(N, K, generator=g) - 0.5).mul_(2 * 448).to(torch.float8_e4m3fn)
    sa = (torch.randn(M, K // 128, generator=g) * 1e-3)
    sb = (torch.randn(K // 128, -(N // -128), generator=g) * 1e-3)
    ad, bd = a.to(DEV), bt.to(DEV).t()
    sad, sbd = sa.t().contiguous().to(DEV).t(), sb.to(DEV)
    out = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)
    out2 = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)
    out4 = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, (sad * 4.0).t().contiguous().t(), sbd, torch.bfloat16)
    torch.cuda.synchronize()
    assert torch.equal(out, out2), "non-deterministic"
    assert torch.equal(out4.float(), out.float() * 4.0), "not linear in a power-of-two scale"
    rows = torch.unique(torch.cat([torch.arange(0, M, max(1, M // 96)), torch.tensor([M - 1])]))
    ref = O.baseline_scaled_mm(a[rows], bt.t(), sa[rows], sb, torch.bfloat16)
    check(out[rows.to(DEV)].cpu(), ref)


@pytest.fixture
def force_stream_k():
    """Stream-K whenever the shape allows it (the cost model keeps small test shapes on whole tiles)."""
    with _ops.stream_k_mode("force"):
        yield


STREAMK = [(1, 2112, 7168), (16, 24576, 7168), (64, 2112, 7168), (130, 4096, 2048), (256, 2112, 7168), (300, 1280, 4096),
           (128, 3200, 1024)]


@pytest.mark.parametrize("M,N,K", STREAMK)
@pytest.mark.parametrize("variant", [-1, 0, 2, 3, 5, 6, 13, 14, 23, 29, 30, 31, 32])
def test_stream_k_matches_whole_tiles_and_oracle(M, N, K, variant, force_stream_k):
    """Stream-K scheduling (tiles cut across CTAs, summed through the workspace by the last arriver)
    gives the oracle's answer, is bit-reproducible, and leaves its counters zeroed for the next launch."""
    skip_if_too_many_tokens(variant, M)
    lib = _lib.load()
    need = lib.sglk_gemm_workspace_bytes(M, N, K, variant, 0, _lib.STREAMK_FORCE)
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M * 7 + N)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.t().contiguous().to(DEV).t(), sb.to(DEV)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    out_sk = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant, 0, True)
    out_sk2 = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant, 0, True)      # reuses the workspace: counters must be back at zero
    out_dp = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant, 0, False)
    torch.cuda.synchronize()
    check(out_sk.cpu(), ref)
    check(out_dp.cpu(), ref)
    assert torch.equal(out_sk, out_sk2), "stream-K result not reproducible"
    if need == 0:
        assert torch.equal(out_sk, out_dp)
    if variant == -1 and M <= 256 and N >= 2048 and lib.sglk_gemm_pick_variant(M, N, K) != 32:
        assert need > 0, "decode shapes on the tile kernels are expected to take the stream-K path"


def test_stream_k_odd_sm_counts(force_stream_k):
    """Unit ranges that start and end in the middle of tiles, for several worker counts."""
    M, N, K = 40, 1664, 2304
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=5)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.t().contiguous().to(DEV).t(), sb.to(DEV)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    for variant in (0, 30, 31, 32):
        for sms in (148, 37, 24, 7, 100):
            out = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant, sms, True)
            torch.cuda.synchronize()
            check(out.cpu(), ref)


@pytest.mark.parametrize("M,N_total,n_lo,n_hi,K", [(300, 1024, 256, 768, 1024), (64, 4608, 2304, 4608, 2048), (1024, 768, 0, 256, 512)])
def test_multi_destination_epilogue(M, N_total, n_lo, n_hi, K):
    """The fused GEMM->all-gather epilogue on ONE GPU: the 'peers' are two more local [M, N_total] buffers.
    Every destination must receive this rank's panel at columns [n_lo, n_hi) and nothing else."""
    N = n_hi - n_lo
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=11)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.t().contiguous().to(DEV).t(), sb.to(DEV)
    bufs = [torch.full((M, N_total), -7.0, dtype=torch.bfloat16, device=DEV) for _ in range(3)]
    _ops._gemm_allgather(ad, bd, sad, sbd, bufs[0], n_lo, [bufs[1].data_ptr(), bufs[2].data_ptr()])
    torch.cuda.synchronize()
    for buf in bufs:
        got = buf.cpu()
        check(got[:, n_lo:n_hi].contiguous(), ref)
        outside = torch.cat([got[:, :n_lo], got[:, n_hi:]], dim=1)
        assert (outside == -7.0).all(), "epilogue wrote outside its panel"
    assert torch.equal(bufs[0], bufs[1]) and torch.equal(bufs[0], bufs[2])


def test_empty_and_degenerate():
    a, b, sa, sb = O.make_gemm_inputs(0, 256, 512, seed=0)
    out = sgl_kernel.fp8_blockwise_scaled_mm(a.to(DEV), b.t().contiguous().to(DEV).t(), sa.to(DEV), sb.to(DEV), torch.bfloat16)
    assert out.shape == (0, 256)


def test_error_behaviour():
    a, b, sa, sb = O.make_gemm_inputs(8, 256, 512, seed=0)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.to(DEV), sb.to(DEV)
    f = sgl_kernel.fp8_blockwise_scaled_mm
    with pytest.raises(RuntimeError, match="column major"):
        f(ad, bd.contiguous(), sad, sbd, torch.bfloat16)
    with pytest.raises(RuntimeError, match="row major"):
        f(ad.t().contiguous().t(), bd, sad, sbd, torch.bfloat16)
    with pytest.raises(RuntimeError, match="Float8"):
        f(ad.view(torch.uint8), bd, sad, sbd, torch.bfloat16)
    with pytest.raises(RuntimeError, match="out_dtype"):
        f(ad, bd, sad, sbd, torch.float32)
    with pytest.raises(RuntimeError, match="scales_a"):
        f(ad, bd, sad[:, :1], sbd, torch.bfloat16)
    with pytest.raises(RuntimeError, match="cannot be multiplied"):
        f(ad[:, :256], bd, sad, sbd, torch.bfloat16)


def test_cuda_graph_capture():
    a, b, sa, sb = O.make_gemm_inputs(64, 512, 1024, seed=7)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.t().contiguous().to(DEV).t(), sb.to(DEV)
    sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)   # warm-up (sets func attributes)
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            out = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)
    g.replay()
    torch.cuda.synchronize()
    check(out.cpu(), O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16))


@pytest.mark.parametrize("M", [300, 1])
def test_host_streamed_pipelines_match_device_result(M):
    """The host-buffer entry points (bench.py's `e2e`): pinned x -> pinned y through the overlapped pipelines equals the
    plain device-side layer call bit for bit, for several consecutive batches (buffer reuse), single rank."""
    from sgl_kernel.sharded import ColumnShardedFp8Linear, HostStreamedShardedFp8Linear

    x, wq_t, ws_t = O.make_linear_inputs(M, 512, 1024, seed=5)
    layer = sgl_kernel.Fp8BlockLinear(wq_t.t().contiguous().to(DEV), ws_t.t().contiguous().to(DEV))
    runners = [sgl_kernel.HostStreamedFp8Linear(layer, 512),
               HostStreamedShardedFp8Linear(ColumnShardedFp8Linear(layer, 512), 512)]
    xs = [(x.float() * (i + 1)).to(torch.bfloat16).pin_memory() for i in range(3)]
    want = [layer(xh.to(DEV)).cpu() for xh in xs]
    for runner in runners:
        ys = [torch.empty(M, 512, dtype=torch.bfloat16).pin_memory() for _ in xs]
        for xh, yh in zip(xs, ys):
            runner.submit(xh, yh)
        runner.synchronize()
        for yh, w in zip(ys, want):
            assert torch.equal(yh, w)


# ---------------------------------------------------------------------------------------------------------------
# decode (swap-AB) kernel: the reference grid's small-M rows (tests/test_fp8_blockwise_gemm.py:88: M in 1, 3, 5, 127,
# 128) on both forms, every token-count bucket boundary, both output dtypes, both scale layouts, ragged N
# ---------------------------------------------------------------------------------------------------------------
DECODE_SHAPES = [(1, 128, 512), (3, 512, 1024), (5, 1024, 4096), (15, 384, 256), (16, 2112, 7168), (17, 256, 1024),
                 (32, 1280, 2048), (33, 640, 512), (64, 4096, 1024), (65, 128, 7168), (127, 4096, 8192), (128, 8192, 512),
                 (129, 512, 1024), (200, 2112, 2048), (255, 384, 4096), (256, 14080, 512), (40, 192, 384), (7, 64, 128)]


@pytest.mark.parametrize("variant", DECODE)
@pytest.mark.parametrize("M,N,K", DECODE_SHAPES)
@pytest.mark.parametrize("sched", ["auto", "force", "never"])
def test_decode_kernel(variant, M, N, K, sched):
    skip_if_too_many_tokens(variant, M)
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=3 * M + N + K)
    odt = torch.bfloat16 if (M + N // 64) % 2 == 0 else torch.float16
    ref = O.baseline_scaled_mm(a, b, sa, sb, odt)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    with _ops.stream_k_mode(sched):
        out_mmaj = _ops._gemm(ad, bd, sa.t().contiguous().to(DEV).t(), sb.to(DEV), odt, variant)          # the reference test's layouts
        out_rowm = _ops._gemm(ad, bd, sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t(), odt, variant)   # the quant op's default
        out_again = _ops._gemm(ad, bd, sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t(), odt, variant)
    torch.cuda.synchronize()
    check(out_mmaj.cpu(), ref)
    assert torch.equal(out_mmaj, out_rowm), "result depends on the scale layout"
    assert torch.equal(out_rowm, out_again), "not reproducible"


def test_decode_is_the_default_for_small_m():
    """the row-balanced swap-AB kernel is the default exactly where it measured faster (gemm.cu pick_variant): up to 64
    tokens; larger batches stay on the token-major tiles"""
    lib = _lib.load()
    for N in (2112, 4096, 7168, 24576, 129280):
        for M in (1, 8, 16, 17, 32, 64):
            assert lib.sglk_gemm_pick_variant(M, N, 7168) == 32
        for M in (65, 128):      # one wave of 256-wide single-CTA tiles where that fills the SMs, else 128-wide + stream-K
            assert lib.sglk_gemm_pick_variant(M, N, 7168) == (14 if N == 24576 else 0)
    assert lib.sglk_gemm_pick_variant(256, 2112, 7168) == 0
    assert lib.sglk_gemm_pick_variant(256, 7168, 7168) == 5          # one wave of 256x128 pair tiles
    assert lib.sglk_gemm_pick_variant(256, 24576, 7168) == 29
    assert lib.sglk_gemm_pick_variant(257, 24576, 7168) == 29
    assert lib.sglk_gemm_pick_variant(16384, 18432, 7168) == 29


@pytest.mark.parametrize("M,N,K", [(16, 2112, 7168), (64, 2112, 7168), (128, 4096, 7168), (5, 1000, 2048), (33, 7168, 1024),
                                   (16, 4096, 7168), (1, 128, 1024), (128, 640, 8192)])
def test_rows_kernel_cluster_split_k(M, N, K):
    """K split over clusters of 1 / 2 / 4 CTAs (partials summed through distributed shared memory in rank order):
    the oracle's answer for every split, reproducible, and the automatic choice is one of them"""
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N + K)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.contiguous().to(DEV), sb.to(DEV)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    outs = {ks: _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, 32, split=ks) for ks in (1, 2, 4)}
    auto = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, 32)
    again = {ks: _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, 32, split=ks) for ks in (2, 4)}
    torch.cuda.synchronize()
    for ks, y in outs.items():
        check(y.cpu(), ref)
    for ks, y in again.items():
        assert torch.equal(y, outs[ks]), f"split {ks} not reproducible"
    assert any(torch.equal(auto, y) for y in outs.values())


def test_rows_kernel_load_forms_agree():
    """one 3-D TMA box per operand per stage (K % 128 == 0) and one 2-D box per K block give the same bits"""
    for M, N, K in [(16, 2112, 7168), (64, 24576, 1024), (5, 1000, 2048), (128, 4096, 1536)]:
        a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N)
        ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
        sad, sbd = sa.contiguous().to(DEV), sb.to(DEV)
        ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
        y3 = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, 32)
        y2 = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, 32, extra_flags=_lib.DECODE_2D_LOADS)
        torch.cuda.synchronize()
        check(y3.cpu(), ref)
        assert torch.equal(y3, y2)


# K tails (K % 128 != 0: TMA zero-fills the last K block, scale index is ceil(K/128)), odd N, unaligned ldc — the
# reference's baseline asserts divisibility (tests/test_fp8_blockwise_gemm.py:48), so the oracle here is its ragged
# extension (torch_oracle._group_broadcast); unpinned by the reference, documented in DESIGN.md
RAGGED = [(37, 130, 144), (300, 1000, 2064), (5, 77, 272), (129, 250, 400), (64, 2111, 1040)]


@pytest.mark.parametrize("M,N,K", RAGGED)
@pytest.mark.parametrize("variant", [-1, 0, 1, 3, 4, 29, 30, 31, 32])
def test_ragged_k_and_n(M, N, K, variant):
    skip_if_too_many_tokens(variant, M)
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N + K)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    out = gpu_mm(a, b, sa, sb, torch.bfloat16, variant)
    check(out, ref)


def test_two_devices_in_one_process():
    """cudaFuncSetAttribute(MaxDynamicSharedMemorySize) belongs to a device's context: the op must work on cuda:1 after
    it has run on cuda:0 in the same process (and back)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    for M, N, K in [(64, 512, 1024), (300, 512, 1024)]:
        a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=1)
        ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
        for dev in ("cuda:0", "cuda:1", "cuda:0"):
            out = sgl_kernel.fp8_blockwise_scaled_mm(a.to(dev), b.t().contiguous().to(dev).t(), sa.t().contiguous().to(dev).t(),
                                                     sb.to(dev), torch.bfloat16)
            torch.cuda.synchronize(dev)
            check(out.cpu(), ref)


def test_capture_without_warmup_is_safe():
    """A first call on a fresh stream INSIDE a graph capture must not cache a workspace whose zero-fill was only
    recorded: the op runs whole tiles for that call, and eager calls / replays afterwards stay correct."""
    a, b, sa, sb = O.make_gemm_inputs(16, 2112, 2048, seed=9)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.t().contiguous().to(DEV).t(), sb.to(DEV)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    st = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            out = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)
        eager = sgl_kernel.fp8_blockwise_scaled_mm(ad, bd, sad, sbd, torch.bfloat16)     # before the first replay
        g.replay()
    torch.cuda.synchronize()
    check(eager.cpu(), ref)
    check(out.cpu(), ref)


def test_graph_keeps_its_workspace_when_the_cache_grows():
    """A graph captured with a stream-K launch points at the workspace of its stream.  When a later eager call needs a
    bigger one, the old buffer must stay alive and zeroed: were it handed back to the allocator, the tensor that gets
    the memory next would put non-zero bytes under the graph's tile counters and the replay would be wrong."""
    a, b, sa, sb = O.make_gemm_inputs(128, 2112, 2048, seed=21)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.to(DEV), sb.to(DEV)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    st = torch.cuda.Stream()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st):
        with _ops.stream_k_mode("force"):
            _ops._gemm(ad, bd, sad, sbd, torch.bfloat16)               # warm-up on the capture stream: creates its workspace
            old = _ops._workspaces[(ad.device.index, st.cuda_stream)]
            with torch.cuda.graph(g, stream=st):
                out = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16)
        grown = _ops._get_workspace(ad.device, st.cuda_stream, old.numel() * 3)
        assert grown.data_ptr() != old.data_ptr() and any(w is old for w in _ops._retired)
        old_ptr = old.data_ptr()
        del old
        torch.cuda.synchronize()
        junk = [torch.full((1 << 20,), 0x5A, dtype=torch.uint8, device=DEV) for _ in range(8)]   # would land on a freed buffer
        assert all(j.data_ptr() != old_ptr for j in junk)
        g.replay()
    torch.cuda.synchronize()
    check(out.cpu(), ref)


def test_reserve_gemm_workspace_gives_side_stream_captures_stream_k():
    """torch.cuda.graph captures on a side stream; a warm-up on the default stream does not create that stream's
    workspace.  reserve_gemm_workspace(stream=...) does, so the captured launch gets one (and stream-K with it)."""
    st = torch.cuda.Stream()
    key = (torch.cuda.current_device(), st.cuda_stream)
    assert key not in _ops._workspaces
    n = sgl_kernel.reserve_gemm_workspace([(128, 2112, 2048), (16384, 4096, 7168)], stream=st)
    assert _ops._workspaces[key].numel() == n >= _lib.load().sglk_gemm_workspace_bytes(16384, 4096, 7168, -1, 0, 0)
    a, b, sa, sb = O.make_gemm_inputs(128, 2112, 2048, seed=22)
    ad, bd, sad, sbd = a.to(DEV), b.t().contiguous().to(DEV).t(), sa.to(DEV), sb.to(DEV)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            with _ops.stream_k_mode("force"):
                out = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16)
            assert _ops._get_workspace(ad.device, st.cuda_stream, 1 << 20) is _ops._workspaces[key]
        with pytest.raises(RuntimeError, match="outside"):
            with torch.cuda.graph(torch.cuda.CUDAGraph(), stream=st):
                sgl_kernel.reserve_gemm_workspace([(1, 1, 16)], stream=st)
    g.replay()
    torch.cuda.synchronize()
    check(out.cpu(), O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16))
