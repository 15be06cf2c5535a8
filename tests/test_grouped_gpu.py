"""GPU parity tests for fp8_blockwise_scaled_grouped_mm (MoE form of the blockwise GEMM) through torch.ops -> C ABI.
Bar: the reference test's own `assert_close(rtol=1e-2, atol=5e-4)` per expert (tests/test_fp8_blockwise_moe.py:164-168)
plus the dense GEMM's bars (rel-err <= 1e-2, cosine >= 0.9999), and bitwise equality with the dense op run per expert."""
import random

import pytest
import torch

from oracle import torch_oracle as O
from test_oracle import GROUPED, load_grouped

import sgl_kernel

pytestmark = pytest.mark.gpu
DEV = "cuda"


def run_grouped(a, bt, sa, sbt, offs, sizes, odt, workspace_bytes=None):
    """a [sum_m,K], bt [E,N,K], sa [sum_m,G], sbt [E,NB,G] (CPU) -> output (CPU), called the way the reference test does."""
    E, N, K = bt.shape
    ad, bd = a.to(DEV), bt.to(DEV).transpose(1, 2)
    sad, sbd = sa.to(DEV), sbt.to(DEV).transpose(1, 2)
    out = torch.full((a.shape[0], N), float("nan"), dtype=odt, device=DEV)
    i64 = lambda: torch.empty(E, dtype=torch.int64, device=DEV)
    lay = lambda: torch.zeros(E, 5, dtype=torch.int32, device=DEV)
    strides = torch.full((E,), K, dtype=torch.int64, device=DEV)
    cstr = torch.full((E,), N, dtype=torch.int64, device=DEV)
    ws = torch.empty(workspace_bytes if workspace_bytes is not None else 1 << 20, dtype=torch.uint8, device=DEV)
    sgl_kernel.fp8_blockwise_scaled_grouped_mm(out, i64(), i64(), i64(), i64(), i64(), ad, bd, sad, sbd, strides, strides,
                                               cstr, lay(), lay(), sizes.to(DEV), offs.to(DEV), ws)
    torch.cuda.synchronize()
    return out.cpu()


def check(out, ref):
    assert not torch.isnan(out.float()).any(), "rows of some expert were not written"
    torch.testing.assert_close(out, ref, rtol=1e-2, atol=5e-4)
    assert O.rel_err(out, ref) <= 1e-2 and O.cosine(out, ref) >= 0.9999


@pytest.mark.parametrize("name", GROUPED)
def test_golden(grouped_golden, name):
    a, bt, sa, sbt, offs, sizes, o, odt = load_grouped(grouped_golden, name)
    check(run_grouped(a, bt, sa, sbt, offs, sizes, odt), o)


def make_case(E, n, k, m_list, seed, odt=torch.bfloat16):
    g = torch.Generator().manual_seed(seed)
    sum_m = sum(m_list)
    a = O.to_fp8_round(torch.randn(sum_m, k, generator=g) * 40)
    bt = O.to_fp8_round(torch.randn(E, n, k, generator=g) * 40)
    sa = torch.randn(sum_m, -(k // -128), generator=g) * 1e-3
    sbt = torch.randn(E, -(n // -128), -(k // -128), generator=g) * 1e-3
    offs = torch.tensor([sum(m_list[:i]) for i in range(E)], dtype=torch.int32)
    sizes = torch.tensor([[m, n, k] for m in m_list], dtype=torch.int32)
    return a, bt, sa, sbt, offs, sizes, odt


@pytest.mark.parametrize("E", [8, 16])
@pytest.mark.parametrize("odt", [torch.half, torch.bfloat16])
def test_reference_parametrisation(E, odt):
    """tests/test_fp8_blockwise_moe.py:61-171: n, k = 16*rand(1..5)*128, m_g = 16*rand(1..64) (smaller n, k draws here
    so the CPU oracle stays in seconds)."""
    random.seed(E)
    n, k = 128 * random.randint(1, 6), 128 * random.randint(1, 8)
    m_list = [16 * random.randint(1, 64) for _ in range(E)]
    a, bt, sa, sbt, offs, sizes, odt = make_case(E, n, k, m_list, E, odt)
    ref = O.grouped_scaled_mm_ref(a, bt.transpose(1, 2), sa, sbt.transpose(1, 2), offs, sizes, odt)
    check(run_grouped(a, bt, sa, sbt, offs, sizes, odt), ref)


@pytest.mark.parametrize("m_list", [[1, 0, 300, 7], [0, 0, 5], [129, 128, 127, 1, 256], [16] * 40, [2000, 3, 0, 0, 90]])
def test_ragged_and_empty_experts(m_list):
    """Experts with 0 rows, 1 row, tile-straddling row counts; N not a multiple of 128; K with 12 blocks and with 5
    (scalar scale path).  Each expert's panel must equal the DENSE op on that expert bit for bit (same tile shape,
    same K order)."""
    E = len(m_list)
    for n, k in [(320, 1536), (256, 640)]:
        a, bt, sa, sbt, offs, sizes, odt = make_case(E, n, k, m_list, sum(m_list) + n)
        out = run_grouped(a, bt, sa, sbt, offs, sizes, odt)
        ref = O.grouped_scaled_mm_ref(a, bt.transpose(1, 2), sa, sbt.transpose(1, 2), offs, sizes, odt)
        check(out, ref)
        for e, m in enumerate(m_list):
            if m == 0:
                continue
            off = int(offs[e])
            dense = sgl_kernel._ops._gemm(a[off:off + m].to(DEV), bt[e].to(DEV).t(), sa[off:off + m].to(DEV),
                                          sbt[e].to(DEV).t(), odt, 1, 0, False)
            assert torch.equal(dense.cpu(), out[off:off + m]), f"expert {e}"


def test_gaps_between_experts_are_untouched():
    """expert_offsets need not be dense: rows that belong to no expert keep their contents."""
    E, n, k = 3, 256, 512
    m_list = [40, 10, 70]
    a, bt, sa, sbt, _, sizes, odt = make_case(E, n, k, [64, 64, 128], 3)
    offs = torch.tensor([0, 64, 128], dtype=torch.int32)
    sizes = torch.tensor([[m, n, k] for m in m_list], dtype=torch.int32)
    out = run_grouped(a, bt, sa, sbt, offs, sizes, odt)
    ref = O.grouped_scaled_mm_ref(a, bt.transpose(1, 2), sa, sbt.transpose(1, 2), offs, sizes, odt)
    for e, m in enumerate(m_list):
        off = int(offs[e])
        torch.testing.assert_close(out[off:off + m], ref[off:off + m], rtol=1e-2, atol=5e-4)
        end = int(offs[e + 1]) if e + 1 < E else out.shape[0]
        assert torch.isnan(out[off + m:end].float()).all()


@pytest.mark.parametrize("m_list,n,k", [([300, 16, 700, 513], 512, 1024), ([257, 255, 256, 1024], 256, 7168),
                                         ([1000, 0, 40], 768, 512)])
def test_pair_tiles_for_prefill_sized_experts(m_list, n, k):
    """>= 256 tokens per expert on average and N % 256 == 0: 256-row CTA-pair tiles whose boxes are stored with the rows
    cut at the expert's end.  Same bits as the dense op per expert; rows between experts keep their contents."""
    E = len(m_list)
    a, bt, sa, sbt, _, _, odt = make_case(E, n, k, [m + 24 for m in m_list], sum(m_list) + k)
    offs = torch.tensor([sum(m + 24 for m in m_list[:i]) for i in range(E)], dtype=torch.int32)     # 24 unused rows after every expert
    sizes = torch.tensor([[m, n, k] for m in m_list], dtype=torch.int32)
    out = run_grouped(a, bt, sa, sbt, offs, sizes, odt)
    for e, m in enumerate(m_list):
        off = int(offs[e])
        assert torch.isnan(out[off + m:off + m + 24].float()).all(), f"rows behind expert {e} were written"
        if m == 0:
            continue
        dense = sgl_kernel._ops._gemm(a[off:off + m].to(DEV), bt[e].to(DEV).t(), sa[off:off + m].to(DEV), sbt[e].to(DEV).t(),
                                      odt, 29 if m > 128 else 1, 0, False)
        assert torch.equal(dense.cpu(), out[off:off + m]), f"expert {e}"
    e0 = O.baseline_scaled_mm(a[:m_list[0]], bt[0].t(), sa[:m_list[0]], sbt[0].t(), odt)
    check(out[:m_list[0]], e0)


def test_errors_and_graph_capture():
    a, bt, sa, sbt, offs, sizes, odt = make_case(4, 256, 512, [32, 64, 16, 48], 9)
    with pytest.raises(RuntimeError, match="workspace too small"):
        run_grouped(a, bt, sa, sbt, offs, sizes, odt, workspace_bytes=32)
    with pytest.raises(RuntimeError, match="Int32"):
        run_grouped(a, bt, sa, sbt, offs.to(torch.int64), sizes, odt)
    # capturable: table build + GEMM, no host sync
    E, N, K = bt.shape
    ad, bd, sad, sbd = a.to(DEV), bt.to(DEV).transpose(1, 2), sa.to(DEV), sbt.to(DEV).transpose(1, 2)
    out = torch.zeros(a.shape[0], N, dtype=odt, device=DEV)
    d = torch.empty(E, dtype=torch.int64, device=DEV)
    lay = torch.zeros(E, 5, dtype=torch.int32, device=DEV)
    ws = torch.empty(1 << 16, dtype=torch.uint8, device=DEV)
    sz, of = sizes.to(DEV), offs.to(DEV)
    call = lambda: sgl_kernel.fp8_blockwise_scaled_grouped_mm(out, d, d, d, d, d, ad, bd, sad, sbd, d, d, d, lay, lay, sz, of, ws)
    call(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            call()
    out.zero_()
    g.replay(); torch.cuda.synchronize()
    ref = O.grouped_scaled_mm_ref(a, bt.transpose(1, 2), sa, sbt.transpose(1, 2), offs, sizes, odt)
    check(out.cpu(), ref)
