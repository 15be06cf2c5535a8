"""Row-sharded quantisation + gated GEMM on ONE GPU (plain-store form of sglk_per_token_group_quant_fp8_bcast): the
'ranks' are consecutive launches into the same local buffers.  The multi-GPU form (NVLS multicast stores, remote
arrival counters) is checked by tools/sharded_quant_check.py under torchrun and by bench.py's matches_replicated flag.
No reference counterpart (the reference has no communication layer): parity is against the repo's own
ops, which are pinned on the oracle elsewhere (test_quant_gpu.py, test_gemm_gpu.py)."""
import pytest
import torch

import sgl_kernel
from oracle import torch_oracle as O
from sgl_kernel import _ops


def _gated(*args, **kw):
    """the gated GEMM with a bounded wait (a few seconds): here a trap is better than a box that hangs"""
    kw.setdefault("wait_polls", 4_000_000)
    return _ops._gemm_gated(*args, **kw)


pytestmark = pytest.mark.gpu
DEV = "cuda"


def make(M, N, K, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g).to(torch.bfloat16)
    wq = ((torch.rand(N, K, generator=g) - 0.5) * 896).to(torch.float8_e4m3fn)
    ws = (torch.rand(-(N // -128), K // 128, generator=g) + 0.5) * 1e-3
    return x.to(DEV), wq.to(DEV), ws.to(DEV)


@pytest.mark.parametrize("M,K", [(256, 1024), (1024, 7168), (512, 128), (2048, 2048)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_quant_bcast_bit_exact_and_counts(M, K, dtype):
    """codes / scales equal the group-quant op's (and therefore the oracle's), every 256-row counter advanced by
    256 * K / 16384 per call"""
    x = torch.randn(M, K, device=DEV).to(dtype)
    q_ref = torch.empty(M, K, dtype=torch.float8_e4m3fn, device=DEV)
    s_ref = torch.empty(M, K // 128, dtype=torch.float32, device=DEV)
    sgl_kernel.sgl_per_token_group_quant_8bit(x, q_ref, s_ref, 128, 1e-10, -448.0, 448.0, enable_v2=False)
    q = torch.zeros(M, K, dtype=torch.uint8, device=DEV)
    s = torch.zeros(M, K // 128, dtype=torch.float32, device=DEV)
    cnt = torch.zeros(M // 256, dtype=torch.int32, device=DEV)
    for rep in range(2):
        half = M // 2 if M >= 512 else M
        for lo in range(0, M, half):                     # two "ranks"
            _ops._quant_bcast(x[lo:lo + half], q.data_ptr() + lo * K, s.data_ptr() + lo * (K // 128) * 4,
                              cnt.data_ptr() + (lo // 256) * 4, multicast=False)
        torch.cuda.synchronize()
        assert torch.equal(q, q_ref.view(torch.uint8))
        assert torch.equal(s, s_ref)
        assert cnt.tolist() == [(rep + 1) * _ops.counter_increments(K)] * (M // 256)
    xq_o, xs_o = O.per_token_group_quant_fp8_ref(x[:64].cpu(), 128)
    assert torch.equal(q[:64].cpu(), xq_o.view(torch.uint8)) and torch.equal(s[:64].cpu(), xs_o)


@pytest.mark.parametrize("M,N,K,world", [(512, 512, 1024, 1), (1024, 768, 2048, 2), (2048, 2304, 1024, 4), (4096, 512, 512, 8),
                                          (2048, 1280, 7168, 8), (1536, 256, 512, 3)])
def test_gated_gemm_matches_plain(M, N, K, world):
    """every panel rotation walks every tile exactly once: same bits as the ungated kernel on the completed buffers"""
    x, wq, ws = make(M, N, K, seed=M + N + K + world)
    layer = sgl_kernel.Fp8BlockLinear(wq, ws)
    xq, xs = layer.quantize_input(x)
    ref = layer.forward_quantized(xq, xs)
    inc = _ops.counter_increments(K)
    cnt = torch.full((M // 256,), 3 * inc, dtype=torch.int32, device=DEV)     # "third call on this buffer"
    for rank in range(world):
        for whole in (False, True):                      # both arrival orders of the other ranks' panels
            out = _gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, cnt, 3 * inc, rank, world, whole_panels=whole)
            torch.cuda.synchronize()
            assert torch.equal(out, ref), f"rank {rank} of {world}, whole_panels={whole}"
    out = _gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, cnt, 2 * inc, 0, world)   # counters ahead of the target
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


def test_quant_then_gated_gemm_stream_order():
    """the pipeline of one rank: its quantisation launches precede the GEMM on the stream; counters accumulate over steps"""
    M, N, K, world = 2048, 1024, 2048, 4
    x, wq, ws = make(M, N, K, seed=5)
    layer = sgl_kernel.Fp8BlockLinear(wq, ws)
    ref = layer(x)
    q = torch.zeros(M, K, dtype=torch.uint8, device=DEV)
    s = torch.zeros(M, K // 128, dtype=torch.float32, device=DEV)
    cnt = torch.zeros(M // 256, dtype=torch.int32, device=DEV)
    rows = M // world
    for step in range(3):
        q.zero_(); s.zero_()
        for r in range(world):
            lo = r * rows
            _ops._quant_bcast(x[lo:lo + rows], q.data_ptr() + lo * K, s.data_ptr() + lo * (K // 128) * 4,
                              cnt.data_ptr() + (lo // 256) * 4, multicast=False)
        out = _gated(q.view(torch.float8_e4m3fn), wq.t(), s, ws.t(), torch.bfloat16, cnt,
                               (step + 1) * _ops.counter_increments(K), step % world, world)
        torch.cuda.synchronize()
        assert torch.equal(out, ref)


def test_gated_errors():
    x, wq, ws = make(512, 512, 1024, seed=1)
    layer = sgl_kernel.Fp8BlockLinear(wq, ws)
    xq, xs = layer.quantize_input(x)
    cnt = torch.zeros(2, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError):        # 512 rows are not 4 whole 256-row panels
        _gated(xq, wq.t(), xs, ws.t(), torch.bfloat16, cnt, 0, 0, 4)
    with pytest.raises(RuntimeError):        # M-major scales cannot be gated
        _gated(xq, wq.t(), xs.t().contiguous().t(), ws.t(), torch.bfloat16, cnt, 0, 0, 1)
    with pytest.raises(RuntimeError):        # rows not a multiple of the counter block
        _ops._quant_bcast(x[:100], xq.data_ptr(), xs.data_ptr(), cnt.data_ptr(), multicast=False)


@pytest.mark.parametrize("M,N,K,world,whole", [(2048, 768, 1024, 4, False), (2048, 512, 2048, 2, True),
                                                 (256 * 66, 256, 512, 2, False)])
def test_gated_gemm_waits_for_rows_that_arrive_during_the_launch(M, N, K, world, whole):
    """The GEMM is launched on buffers that hold poison (NaN codes, NaN scales) and counters below the target; a side
    stream delivers rows and counters block by block WHILE it runs — late, and not in the order the kernel walks them —
    the way the copy-engine transport of forward_sharded_quant does (data first, then the counter, in stream order).
    Exercises the polling loop, the publication of the arrival mask to the promotion warps, the proxy fence after a late
    arrival and the late scale load; the last case has more than 64 m-blocks (one counter polled per tile).  A tile
    started before its rows were there would put NaN into the output."""
    x, wq, ws = make(M, N, K, seed=M + K + world)
    layer = sgl_kernel.Fp8BlockLinear(wq, ws)
    xq, xs = layer.quantize_input(x)
    ref = layer.forward_quantized(xq, xs)
    inc = _ops.counter_increments(K)
    nblk = M // 256
    done = torch.full((nblk,), 7 * inc, dtype=torch.int32, device=DEV)        # "seventh call on this buffer"
    q = torch.full((M, K), 0x7F, dtype=torch.uint8, device=DEV)               # 0x7F = e4m3 NaN
    s = torch.full((M, K // 128), float("nan"), dtype=torch.float32, device=DEV)
    cnt = torch.full((nblk,), 6 * inc, dtype=torch.int32, device=DEV)
    xq8 = xq.view(torch.uint8)
    side = torch.cuda.Stream()
    torch.cuda.synchronize()
    step = max(1, nblk // 16)
    with torch.cuda.stream(side):
        torch.cuda._sleep(3_000_000)                                          # ~1.5 ms: the GEMM is running and waiting by then
        for b0 in reversed(range(0, nblk, step)):                             # last blocks first: the walk's first tiles wait longest
            r0, r1 = b0 * 256, min(nblk, b0 + step) * 256
            q[r0:r1].copy_(xq8[r0:r1], non_blocking=True)
            s[r0:r1].copy_(xs[r0:r1], non_blocking=True)
            cnt[b0:b0 + step].copy_(done[b0:b0 + step], non_blocking=True)
            torch.cuda._sleep(100_000)
    # 128 of the 148 SMs: the delivery needs none for its copies, but the spinning helper kernels must stay schedulable
    out = _gated(q.view(torch.float8_e4m3fn), wq.t(), s, ws.t(), torch.bfloat16, cnt, 7 * inc, world - 1, world,
                 num_sms=128, whole_panels=whole)
    torch.cuda.synchronize()
    assert not torch.isnan(out.float()).any(), "a tile was started before its rows had arrived"
    assert torch.equal(out, ref)
