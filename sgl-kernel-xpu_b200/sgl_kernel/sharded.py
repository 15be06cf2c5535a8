"""Column-(N-)sharded FP8 block linear over the GPUs of one NVSwitch box.

One process per GPU.  Rank r owns weight rows [r*N/g, (r+1)*N/g) — a multiple of 128, so no 128x128
scale block straddles ranks — and the matching rows of the weight-scale matrix.  Activations are
replicated and quantised on every rank (cheap, HBM-bound); each rank's GEMM produces its [M, N/g]
column panel with no reduction.  Only when the consumer needs the full [M, N] is one all-gather issued
(NCCL over NVLink; gloo in the CPU tests).  The reference has no communication layer at all
(SURVEY.md §2), so there is no reference file to cite for this part; the partitioning follows
BASELINE.json's north_star.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int, block: int = 128) -> Tuple[int, int]:
    """Rows [lo, hi) of the [N, K] weight owned by `rank`: whole 128-row blocks, spread as evenly as
    possible (earlier ranks take the remainder); the last block may be ragged (N % 128 != 0)."""
    blocks = -(n // -block)
    base, rem = divmod(blocks, world)
    lo_b = rank * base + min(rank, rem)
    hi_b = lo_b + base + (1 if rank < rem else 0)
    return min(lo_b * block, n), min(hi_b * block, n)


def shard_weight(weight_q: torch.Tensor, weight_scale: torch.Tensor, world: int, rank: int):
    """Slice a full [N,K] code matrix and its [N/128,K/128] scales down to this rank's panel."""
    lo, hi = shard_bounds(weight_q.shape[0], world, rank)
    return weight_q[lo:hi].contiguous(), weight_scale[lo // 128 : -(hi // -128)].contiguous()


class ColumnShardedFp8Linear:
    def __init__(self, local_linear, n_total: int, group: Optional[dist.ProcessGroup] = None):
        self.local = local_linear
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n_total = n_total
        self.bounds = [shard_bounds(n_total, self.world, r) for r in range(self.world)]
        assert self.bounds[self.rank][1] - self.bounds[self.rank][0] == local_linear.out_features

    def forward_local(self, x: torch.Tensor) -> torch.Tensor:
        return self.local(x)

    def gather(self, y_local: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """All-gather the column panels into [M, N]: one all_gather_into_tensor (rank-major
        [g, M, N/g]) and one strided copy into the column-major-of-panels [M, N] result."""
        if self.world == 1:
            return y_local
        m = y_local.shape[0]
        widths = [hi - lo for lo, hi in self.bounds]
        if out is None:
            out = torch.empty((m, self.n_total), dtype=y_local.dtype, device=y_local.device)
        wmax = max(widths)
        staged = torch.empty((self.world * m, wmax), dtype=y_local.dtype, device=y_local.device)
        if len(set(widths)) == 1:
            dist.all_gather_into_tensor(staged, y_local.contiguous(), group=self.group)
            out.view(m, self.world, wmax).copy_(staged.view(self.world, m, wmax).permute(1, 0, 2))
        else:
            # ragged last panel (N % (128*g) != 0): pad to the widest panel, gather, slice
            mine = torch.zeros((m, wmax), dtype=y_local.dtype, device=y_local.device)
            mine[:, : y_local.shape[1]] = y_local
            dist.all_gather_into_tensor(staged, mine, group=self.group)
            for r, (lo, hi) in enumerate(self.bounds):
                out[:, lo:hi] = staged[r * m : (r + 1) * m, : hi - lo]
        return out

    # ---- fused GEMM -> all-gather over NVLink peer memory -------------------------------------------
    def enable_fused_gather(self, max_tokens: int, dtype: torch.dtype = torch.bfloat16) -> None:
        """Allocate the symmetric [max_tokens, N] output buffer and exchange peer mappings (collective)."""
        import torch.distributed._symmetric_memory as symm_mem

        dev = self.local.weight_q.device
        self._sym_out = symm_mem.empty((max_tokens, self.n_total), dtype=dtype, device=dev)
        self._hdl = symm_mem.rendezvous(self._sym_out, self.group if self.group is not None else dist.group.WORLD)
        # peers in ring order starting after this rank, so that at any instant the ranks target different receivers
        self._peer_ptrs = [int(self._hdl.buffer_ptrs[(self.rank + i) % self.world]) for i in range(1, self.world)]

    def forward_fused_gather(self, x: torch.Tensor, multicast: Optional[bool] = None) -> torch.Tensor:
        """[M, N] on every rank: each rank's GEMM epilogue TMA-stores its output boxes into its own and all
        peers' buffers while the MMAs still run; two symmetric-memory barriers bracket the step (peers done
        reading the previous result / all panels landed).  With `multicast=True` (and NVLS available) the
        epilogue uses multimem.st on the multicast mapping instead of one store per peer.
        Returns a view of the symmetric buffer."""
        from . import _ops

        m = x.shape[0]
        out = self._sym_out[:m]
        x_q, x_s = self.local.quantize_input(x)
        self._hdl.barrier(channel=0)
        lo, _ = self.bounds[self.rank]
        if multicast is None:   # measured: multicast wins from 4 GPUs up (8 GPUs: 0.98 vs 1.18 ms), unicast at 2
            multicast = self.world >= 4
        mc = int(self._hdl.multicast_ptr) if (multicast and self._hdl.multicast_ptr) else 0
        _ops._gemm_allgather(x_q, self.local.weight_q.t(), x_s, self.local.weight_scale.t(), out, lo, self._peer_ptrs,
                             multicast_ptr=mc)
        self._hdl.barrier(channel=1)
        return out

    def forward(self, x: torch.Tensor, gather_output: bool = True) -> torch.Tensor:
        y = self.forward_local(x)
        return self.gather(y) if gather_output else y

    __call__ = forward
