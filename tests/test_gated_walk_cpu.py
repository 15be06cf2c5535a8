"""CPU suite: host-side bookkeeping of the row-sharded quantisation + gated GEMM.  No device is touched:
sglk_gated_walk_order is the host copy of the kernel's tile walk (csrc/gemm.cu tile_coords_panels), and the counter
arithmetic of sgl_kernel/sharded.py forward_sharded_quant is replayed on a world_size-2 gloo group."""
import ctypes
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sgl_kernel import _lib, _ops


def walk(M, N, K, rank, world, whole):
    lib = _lib.load()
    n = lib.sglk_gated_walk_order(None, M, N, K, rank, world, int(whole))
    assert n > 0
    buf = (ctypes.c_int32 * (2 * n))()
    assert lib.sglk_gated_walk_order(ctypes.cast(buf, ctypes.c_void_p), M, N, K, rank, world, int(whole)) == n
    return [(buf[2 * t], buf[2 * t + 1]) for t in range(n)]


@pytest.mark.parametrize("M,N,K,world", [(16384, 2304, 7168, 8), (16384, 4608, 7168, 4), (16384, 9216, 7168, 2), (4096, 512, 512, 8),
                                          (1536, 256, 1024, 3), (512, 1024, 2048, 1), (32768, 2304, 7168, 8)])
@pytest.mark.parametrize("whole", [False, True])
def test_walk_is_a_permutation_in_arrival_order(M, N, K, world, whole):
    num_m, num_n = M // 256, -(N // -256)
    own = num_m // world
    for rank in range(world):
        tiles = walk(M, N, K, rank, world, whole)
        assert sorted(tiles) == [(m, n) for m in range(num_m) for n in range(num_n)], "every tile exactly once"
        # own panel first
        first = tiles[: own * num_n]
        assert {m for m, _ in first} == set(range(rank * own, (rank + 1) * own))
        # the order in which m-blocks are first touched afterwards is the arrival order
        seen, order = set(), []
        for m, _ in tiles[own * num_n:]:
            if m not in seen:
                seen.add(m)
                order.append(m)
        peers = [(rank + 1 + i) % world for i in range(world - 1)]
        if whole:      # one panel after the other, nearest rank first; blocks of a panel stay together
            assert [m // own for m in order[::own]] == peers
            for i, pnl in enumerate(peers):
                assert sorted(order[i * own:(i + 1) * own]) == list(range(pnl * own, (pnl + 1) * own))
        elif world > 1:    # slice by slice: the j-th block of every other panel before any (j+g)-th, g = slices per group
            w1 = world - 1
            slices_per_group = max(1, 8 // w1)
            pos = {m: i for i, m in enumerate(order)}
            for m in order:
                j = m % own
                for m2 in order:
                    if (m2 % own) // slices_per_group > j // slices_per_group:
                        assert pos[m2] > pos[m]


def test_walk_rejects_what_the_kernel_rejects():
    lib = _lib.load()
    assert lib.sglk_gated_walk_order(None, 512, 512, 1024, 0, 4, 0) == -1      # 512 rows: not 4 whole 256-row panels
    assert lib.sglk_gated_walk_order(None, 1024, 128, 1024, 0, 2, 0) == -1     # N < 256
    assert lib.sglk_gated_walk_order(None, 1024, 512, 1024, 2, 2, 0) == -1     # rank out of range


def test_counter_increments():
    assert _ops.counter_increments(7168) == 112 and _ops.counter_increments(128) == 2
    for k in (128, 1024, 7168, 16384):
        assert _ops.ROWS_PER_COUNTER * k % 16384 == 0, "a 256-row block is a whole number of 1024-vector chunks"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _bookkeeping_worker(rank, world, port, q):
    """Replays forward_sharded_quant's counter protocol with CPU tensors: two buffer sets, each rank adds
    counter_increments(K) to its own blocks and 'sends' them (all_gather stands in for the copies); the target the GEMM
    would wait for must be reached by every block after every step, and never before the sending step."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    M, K = 2048, 7168
    inc, blocks, own = _ops.counter_increments(K), M // 256, M // 256 // world
    counters = [torch.zeros(blocks, dtype=torch.int64) for _ in range(2)]
    ok = True
    for step in range(7):
        b, calls = step & 1, step // 2 + 1
        target = calls * inc
        ok = ok and bool((counters[b] < target).all())              # nobody can see this step's rows before they are sent
        mine = counters[b][rank * own:(rank + 1) * own] + inc        # the local quant kernel's increments
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)                              # the copies that follow the data
        counters[b] = torch.cat(gathered)
        ok = ok and bool((counters[b] - target >= 0).all()) and bool((counters[b] == target).all())
    q.put((rank, ok))
    dist.destroy_process_group()


def test_counter_protocol_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_bookkeeping_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == {0: True, 1: True}
