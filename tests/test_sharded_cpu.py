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
