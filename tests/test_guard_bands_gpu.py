"""Out-of-bounds and race checks that do not need compute-sanitizer (closed on this GPU pool, see
profiles/sanitizer_r02.txt):
  * guard bands — every output and every workspace is a view into a larger allocation whose 4 KiB before and after the
    view hold a canary pattern; any store outside the declared extent (a predication bug in a ragged tile, a workspace
    under-reported by sglk_gemm_workspace_bytes, a TMA store box past the tensor) changes it;
  * exact workspace — the GEMM is called through the C ABI with a workspace of exactly the bytes the library asks for;
  * repeatability — kernels whose warps hand work to each other through shared memory, TMEM and global counters
    (stream-K fix-up, swap-AB decode, fused quantisation epilogue, grouped GEMM) must return bit-identical results on
    every one of 25 runs; a missing barrier or fence shows up as run-to-run differences."""
import os

import pytest
import torch

from oracle import torch_oracle as O

import sgl_kernel
from sgl_kernel import _lib, _ops

pytestmark = pytest.mark.gpu
DEV = "cuda"
GUARD = 4096
CANARY = 0xA5


class Guarded:
    """A tensor view of `nbytes` at `offset` bytes into a canary-filled allocation."""

    def __init__(self, nbytes, misalign=0):
        self.nbytes = nbytes
        self.off = GUARD + misalign
        self.raw = torch.full((GUARD + misalign + nbytes + GUARD,), CANARY, dtype=torch.uint8, device=DEV)

    def view(self, shape, dtype, strides=None):
        flat = self.raw[self.off:self.off + self.nbytes].view(dtype)
        return flat.view(shape) if strides is None else flat.as_strided(shape, strides)

    def data(self):
        return self.raw[self.off:self.off + self.nbytes]

    def assert_intact(self, what):
        before = self.raw[:self.off]
        after = self.raw[self.off + self.nbytes:]
        nb, na = int((before != CANARY).sum()), int((after != CANARY).sum())
        assert nb == 0 and na == 0, f"{what}: {nb} bytes before and {na} bytes after the buffer were overwritten"


QUANT_SHAPES = [(1, 128), (5, 2048), (130, 7168), (33, 384), (257, 1024)]


@pytest.mark.parametrize("m,k", QUANT_SHAPES)
@pytest.mark.parametrize("colmajor", [False, True])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_group_quant_stays_inside_its_outputs(m, k, colmajor, dtype):
    x = (torch.randn(m, k) * 3).to(dtype).to(DEV)
    g = k // 128
    gq, gs = Guarded(m * k), Guarded(m * g * 4)
    q = gq.view((m, k), torch.float8_e4m3fn)
    s = gs.view((m, g), torch.float32, (1, m) if colmajor else None)
    sgl_kernel.sgl_per_token_group_quant_8bit(x, q, s, 128, 1e-10, -448.0, 448.0, enable_v2=False)
    torch.cuda.synchronize()
    gq.assert_intact("quant codes")
    gs.assert_intact("quant scales")
    q_ref, s_ref = O.per_token_group_quant_fp8_ref(x.cpu(), 128)
    assert torch.equal(q.cpu().view(torch.uint8), q_ref.view(torch.uint8)) and torch.equal(s.cpu(), s_ref)


@pytest.mark.parametrize("tokens,hidden", [(1, 4), (3, 1076), (7, 7168), (130, 16392)])
def test_token_quant_stays_inside_its_outputs(tokens, hidden):
    x = torch.randn(tokens, hidden, dtype=torch.bfloat16, device=DEV)
    gq, gs = Guarded(tokens * hidden), Guarded(tokens * 4)
    q = gq.view((tokens, hidden), torch.float8_e4m3fn)
    s = gs.view((tokens, 1), torch.float32)
    sgl_kernel.sgl_per_token_quant_fp8(x, q, s)
    torch.cuda.synchronize()
    gq.assert_intact("token quant codes")
    gs.assert_intact("token quant scales")


def gemm_guarded(a, b, sa, sb, odt, variant, mode, ldc=None):
    """fp8_blockwise_scaled_mm through the C ABI: guarded output (optionally with a padded leading dimension) and a
    guarded workspace of exactly sglk_gemm_workspace_bytes."""
    lib = _lib.load()
    M, K = a.shape
    N = b.shape[1]
    ldc = ldc or N
    go = Guarded(((M - 1) * ldc + N) * 2)
    out = go.view((M, N), odt, (ldc, 1))
    flags = {"auto": _lib.STREAMK_AUTO, "force": _lib.STREAMK_FORCE, "never": _lib.STREAMK_NEVER}[mode]
    need = lib.sglk_gemm_workspace_bytes(M, N, K, variant, 0, flags)
    gw = Guarded(max(need, 16))
    gw.data().zero_()
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.sglk_fp8_blockwise_scaled_mm_ex(
        a.data_ptr(), b.data_ptr(), sa.data_ptr(), sb.data_ptr(), out.data_ptr(), M, N, K,
        a.stride(0) if M > 1 else max(K, 16), b.stride(1) if N > 1 else max(K, 16), ldc, sa.stride(0), sa.stride(1),
        sb.stride(0), sb.stride(1), _ops._GEMM_OUT[odt], variant, 0, flags, gw.data().data_ptr() if need else None,
        need, stream)
    _lib.check(rc, "fp8_blockwise_scaled_mm")
    torch.cuda.synchronize()
    go.assert_intact(f"GEMM output (variant {variant}, {mode})")
    gw.assert_intact(f"GEMM workspace (variant {variant}, {mode}, {need} bytes)")
    if need:     # the kernel leaves its tile counters zeroed for the next launch
        assert int(gw.data()[: min(need, 1 << 20)].view(torch.int32).abs().sum()) == 0, "tile counters not reset"
    return out


GEMM_SHAPES = [(1, 128, 512), (5, 2112, 1024), (37, 130, 144), (129, 250, 400), (130, 384, 7168), (300, 1000, 2064),
               (256, 640, 2048)]


@pytest.mark.parametrize("variant", [-1, 0, 1, 2, 3, 4, 5, 6, 13, 14, 23, 29, 30, 31])
@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("mode", ["auto", "force"])
def test_gemm_stays_inside_output_and_workspace(variant, M, N, K, mode):
    if variant in (30, 31) and M > 256:
        pytest.skip("decode kernel: M <= 256")
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M * 7 + N)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t()
    out = gemm_guarded(ad, bd, sad, sbd, torch.bfloat16, variant, mode)
    assert O.rel_err(out.cpu(), ref) <= 1e-2 and O.cosine(out.cpu(), ref) >= 0.9999
    # padded rows: the gap between rows belongs to somebody else too
    M2 = min(M, 40)
    ldc = N + 24 if N % 8 == 0 else N + 3
    go = gemm_guarded(ad[:M2], bd, sad[:M2], sbd, torch.bfloat16, variant, mode, ldc=ldc)
    assert O.rel_err(go.cpu(), ref[:M2]) <= 1e-2


def test_gemm_row_gaps_are_untouched():
    """output with ldc > N: the bytes between the rows keep their canary"""
    M, N, K = 200, 384, 1024
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=5)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t()
    for variant in (0, 1, 3, 29, 30):
        ldc = N + 64
        out = gemm_guarded(ad, bd, sad, sbd, torch.bfloat16, variant, "auto", ldc=ldc)
        raw = out.as_strided((M - 1, ldc - N), (ldc, 1), out.storage_offset() + N).view(torch.uint8)
        assert int((raw != CANARY).sum()) == 0, f"variant {variant} wrote between the rows"


@pytest.mark.parametrize("M,N,K", [(5, 256, 512), (130, 384, 256), (300, 1024, 512)])
@pytest.mark.parametrize("fuse", [False, True])
def test_fused_quant_epilogue_is_repeatable(M, N, K, fuse):
    if fuse and N % 256:
        pytest.skip("gate|up mode: N % 256 == 0")
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=M + N)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.contiguous().to(DEV) * 30.0, sb.t().contiguous().to(DEV).t()
    if fuse:
        from sgl_kernel import _lib
        if "diag" not in os.path.basename(_lib.LIB_PATH):
            pytest.skip("silu(gate)*up epilogue: laboratory build only")
    q0, s0 = sgl_kernel.fp8_blockwise_scaled_mm_quant(ad, bd, sad, sbd, torch.bfloat16, fuse_silu_and_mul=fuse)
    for _ in range(24):
        q, s = sgl_kernel.fp8_blockwise_scaled_mm_quant(ad, bd, sad, sbd, torch.bfloat16, fuse_silu_and_mul=fuse)
        assert torch.equal(q.view(torch.uint8), q0.view(torch.uint8)) and torch.equal(s, s0)


@pytest.mark.parametrize("variant,M,N,K", [(0, 64, 2112, 7168), (3, 300, 1280, 4096), (13, 384, 1280, 7168), (29, 384, 1280, 7168),
                                          (30, 16, 2112, 7168), (31, 16, 2112, 7168), (30, 200, 2112, 2048), (31, 64, 4096, 7168),
                                          (2, 128, 2304, 4096), (23, 256, 1024, 2048)])
def test_stream_k_fixup_is_repeatable(variant, M, N, K):
    """deterministic fix-up order: the last arriver adds the partials in worker order, whoever it is"""
    a, b, sa, sb = O.make_gemm_inputs(M, N, K, seed=variant + M)
    ad, bd = a.to(DEV), b.t().contiguous().to(DEV).t()
    sad, sbd = sa.contiguous().to(DEV), sb.t().contiguous().to(DEV).t()
    with _ops.stream_k_mode("force"):
        first = _ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant)
        for _ in range(24):
            assert torch.equal(_ops._gemm(ad, bd, sad, sbd, torch.bfloat16, variant), first)
    ref = O.baseline_scaled_mm(a, b, sa, sb, torch.bfloat16)
    assert O.rel_err(first.cpu(), ref) <= 1e-2


def test_grouped_gemm_stays_inside_output_and_is_repeatable():
    from test_grouped_gpu import make_case

    m_list = [1, 0, 300, 7, 129, 64]
    E, n, k = len(m_list), 512, 1024
    a, bt, sa, sbt, offs, sizes, odt = make_case(E, n, k, m_list, seed=3)
    ad, bd = a.to(DEV), bt.to(DEV).transpose(1, 2)
    sad, sbd = sa.to(DEV), sbt.to(DEV).transpose(1, 2)
    i64 = lambda: torch.empty(E, dtype=torch.int64, device=DEV)
    lay = lambda: torch.zeros(E, 5, dtype=torch.int32, device=DEV)
    strides = torch.full((E,), k, dtype=torch.int64, device=DEV)
    cstr = torch.full((E,), n, dtype=torch.int64, device=DEV)
    first = None
    for _ in range(10):
        go, gw = Guarded(sum(m_list) * n * 2), Guarded(1 << 20)
        out = go.view((sum(m_list), n), odt)
        ws = gw.view((1 << 20,), torch.uint8)
        sgl_kernel.fp8_blockwise_scaled_grouped_mm(out, i64(), i64(), i64(), i64(), i64(), ad, bd, sad, sbd, strides,
                                                   strides, cstr, lay(), lay(), sizes.to(DEV), offs.to(DEV), ws)
        torch.cuda.synchronize()
        go.assert_intact("grouped GEMM output")
        gw.assert_intact("grouped GEMM workspace")
        if first is None:
            first = out.clone()
        assert torch.equal(out, first)


@pytest.mark.parametrize("M,N,K,world", [(1024, 768, 1024, 2), (2048, 1280, 2048, 4)])
def test_row_sharded_quant_and_gated_gemm_stay_inside_their_buffers_and_repeat(M, N, K, world):
    """sglk_per_token_group_quant_fp8_bcast writes codes, scales and arrival counters at raw addresses;
    sglk_fp8_blockwise_scaled_mm_gated polls the counters and hands an arrival mask between warps through shared memory.
    All four buffers are guarded views; 10 runs (counters accumulating, each rank's panel walk in turn) must leave the
    canaries alone and give the same bits every time."""
    lib = _lib.load()
    g = K // 128
    x = torch.randn(M, K, device=DEV).to(torch.bfloat16)
    wq = ((torch.rand(N, K, device=DEV) - 0.5) * 896).to(torch.float8_e4m3fn)
    ws = (torch.rand(-(N // -128), g, device=DEV) + 0.5) * 1e-3
    gq, gs, gc = Guarded(M * K), Guarded(M * g * 4), Guarded((M // 256) * 4)
    gc.data().zero_()
    q, s, cnt = gq.view((M, K), torch.uint8), gs.view((M, g), torch.float32), gc.view((M // 256,), torch.int32)
    rows = M // world
    first = None
    for run in range(10):
        for r in range(world):
            lo = r * rows
            _ops._quant_bcast(x[lo:lo + rows], q.data_ptr() + lo * K, s.data_ptr() + lo * g * 4,
                              cnt.data_ptr() + (lo // 256) * 4, multicast=False)
        go = Guarded(M * N * 2)
        out = go.view((M, N), torch.bfloat16)
        rc = lib.sglk_fp8_blockwise_scaled_mm_gated(
            q.data_ptr(), wq.data_ptr(), s.data_ptr(), ws.data_ptr(), out.data_ptr(), M, N, K, K, K, N, g, 1, 1, g,
            _lib.BF16, 0, cnt.data_ptr(), (run + 1) * _ops.counter_increments(K), run % world, world, run & 1, 4_000_000,
            torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.sglk_error_string(rc)
        torch.cuda.synchronize()
        for name, gb in (("codes", gq), ("scales", gs), ("counters", gc), ("gated GEMM output", go)):
            gb.assert_intact(name)
        assert cnt.tolist() == [(run + 1) * _ops.counter_increments(K)] * (M // 256)
        if first is None:
            first = out.clone()
            ref = sgl_kernel.Fp8BlockLinear(wq, ws)(x)
            assert torch.equal(first, ref)
        assert torch.equal(out, first), f"run {run} differs from run 0"
