"""CPU suite, part 3: the N>1 host logic (partitioning + all-gather) on world_size-2 and -3 `gloo`
process groups.  The per-rank GEMM is stood in by the CPU oracle (allowed in tests only): what is
under test is that the column panels and their scale rows are cut on 128-block boundaries and that
the gathered [M, N] equals the unsharded result."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_oracle as O
from sgl_kernel.sharded import ColumnShardedFp8Linear, shard_bounds, shard_weight


@pytest.mark.parametrize("n,world", [(18432, 1), (18432, 2), (18432, 4), (18432, 8), (2112, 4), (28672, 8), (300, 3)])
def test_shard_bounds_cover_n_on_block_boundaries(n, world):
    prev = 0
    for r in range(world):
        lo, hi = shard_bounds(n, world, r)
        assert lo == prev and lo % 128 == 0 and (hi % 128 == 0 or hi == n) and hi >= lo
        prev = hi
    assert prev == n
    if n % (128 * world) == 0:
        assert len({shard_bounds(n, world, r)[1] - shard_bounds(n, world, r)[0] for r in range(world)}) == 1


class _OracleLinear:
    """CPU stand-in with the Fp8BlockLinear interface."""

    def __init__(self, wq, ws):
        self.wq, self.ws = wq, ws
        self.out_features, self.in_features = wq.shape

    def __call__(self, x):
        return O.fp8_linear_ref(x, self.wq.t(), self.ws.t())


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        M, K = 6, 256
        x = torch.randn(M, K).to(torch.bfloat16)
        w = torch.randn(n, K)
        wq, ws = O.per_block_cast_to_fp8(w)
        lq, ls = shard_weight(wq, ws, world, rank)
        layer = ColumnShardedFp8Linear(_OracleLinear(lq, ls), n)
        y = layer(x, gather_output=True)
        full = O.fp8_linear_ref(x, wq.t(), ws.t())
        ok = O.rel_err(y, full) < 2e-3 and y.shape == (M, n)
        lo, hi = shard_bounds(n, world, rank)
        ok = ok and torch.equal(layer(x, gather_output=False), y[:, lo:hi])
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 512), (2, 320), (3, 640)])
def test_column_sharded_linear_gloo(world, n):
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, n, ret), nprocs=world, join=True)
        assert dict(ret) == {r: True for r in range(world)}


# ---- row-chunked upload path of HostStreamedShardedFp8Linear: quantise own rows, gather codes + scales ----
from sgl_kernel.sharded import gather_row_chunks, row_chunk  # noqa: E402


@pytest.mark.parametrize("m,world", [(16384, 8), (6, 2), (7, 3), (1, 4), (0, 2)])
def test_row_chunk_covers_m(m, world):
    c = row_chunk(m, world)
    assert c * world >= m and (m == 0 or (c - 1) * world < m)
    covered = sum(max(0, min((r + 1) * c, m) - min(r * c, m)) for r in range(world))
    assert covered == m


def _row_worker(rank, world, port, m, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(3)
        K = 256
        x = torch.randn(m, K).to(torch.bfloat16)                  # the same batch on every rank
        q_full, s_full = O.per_token_group_quant_fp8_ref(x, 128)
        c = row_chunk(m, world)
        xq = torch.zeros(c * world, K, dtype=torch.float8_e4m3fn)
        xs = torch.zeros(c * world, K // 128, dtype=torch.float32)
        lo, hi = min(rank * c, m), min(rank * c + c, m)
        if hi > lo:
            q, s = O.per_token_group_quant_fp8_ref(x[lo:hi], 128)
            xq[lo:hi] = q
            xs[lo:hi] = s
        gather_row_chunks(xq, c, rank)
        gather_row_chunks(xs, c, rank)
        ok = torch.equal(xq[:m].view(torch.uint8), q_full.view(torch.uint8)) and torch.equal(xs[:m], s_full)
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,m", [(2, 6), (3, 7), (2, 1)])
def test_row_chunk_gather_equals_full_quant_gloo(world, m):
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_row_worker, args=(world, port, m, ret), nprocs=world, join=True)
        assert dict(ret) == {r: True for r in range(world)}
