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

    # ---- row-sharded activation quantisation, broadcast over NVLink while the GEMM runs ------------------------
    def sharded_quant_supported(self, tokens: int) -> bool:
        """whole 256-row blocks per rank (one arrival counter per block of the paired GEMM tiles)"""
        return self.world >= 1 and tokens > 0 and tokens % (256 * self.world) == 0 and self.local.out_features >= 256

    def enable_sharded_quant(self, max_tokens: int) -> bool:
        """Allocate the symmetric activation buffers (two sets: codes [M, K] e4m3, scales [M, K/128] fp32, one arrival
        counter per 256 rows) and exchange the mappings (collective)."""
        import torch.distributed._symmetric_memory as symm_mem

        dev = self.local.weight_q.device
        k = self.local.in_features
        assert k % 128 == 0 and max_tokens % 256 == 0
        al = lambda v: -(-v // 256) * 256
        self._sq_codes, self._sq_scales, self._sq_cnt = al(max_tokens * k), al(max_tokens * (k // 128) * 4), al((max_tokens // 256) * 4)
        per_set = self._sq_codes + self._sq_scales + self._sq_cnt
        self._sq_buf = symm_mem.empty((2 * per_set,), dtype=torch.uint8, device=dev)
        self._sq_buf.zero_()
        self._sq_hdl = symm_mem.rendezvous(self._sq_buf, self.group if self.group is not None else dist.group.WORLD)
        self._sq_per_set, self._sq_tokens, self._sq_step = per_set, max_tokens, 0
        # how long the gated GEMM waits for a late rank: 0 = as long as it takes (the default, like any collective);
        # n > 0 = trap after n polling rounds of ~1-2 us — a STICKY failure that costs this process its CUDA context, so
        # only for runs where a hang is worse (benchmarks and tests on boxes with a wall-clock limit)
        if not hasattr(self, "gated_wait_polls"):
            self.gated_wait_polls = 0
        self._sq_hdl.barrier(channel=0)                      # everybody's counters are zero before anybody adds to them
        mc = getattr(self._sq_hdl, "multicast_ptr", 0)
        self._sq_mc = int(mc) if mc else 0
        # copy-engine transport: every peer's buffer as a tensor of this process, a side stream for the sends
        self._sq_peer = [self._sq_buf if r == self.rank else self._sq_hdl.get_buffer(r, (2 * per_set,), torch.uint8, 0)
                         for r in range(self.world)]
        self._sq_copy_stream = torch.cuda.Stream(dev)
        self._sq_ev = torch.cuda.Event()
        return True

    def forward_sharded_quant(self, x: torch.Tensor, transport: str = "copy") -> torch.Tensor:
        """This rank's [M, N/g] panel from REPLICATED bf16 activations `x` [M, K] — like forward_local, but rank r
        quantises only rows [r, r+1) * M/g; the codes, scales and per-256-row arrival counters reach every other
        rank's buffers over NVLink while the GEMM (sglk_fp8_blockwise_scaled_mm_gated) already runs: it starts on its own
        rows at once and on the others' as their counters arrive.  Bit-identical to forward_local (quantisation is
        row-local).  Collective: every rank calls it with the same batch, in the same order.

        transport "copy" (default): the quant kernel writes locally; the copy engines then push the panel to the peers
          on a side stream, nearest consumer first (rank r-1 walks panel r right after its own, rank r-2 after that,
          ...), each peer's data followed by its counters — at any time every rank sends to one peer and receives from
          one, and no SM is involved, so the sends overlap the GEMM completely;
        transport "multicast": the quant kernel itself stores to the NVLS multicast mapping (multimem.st) — one pass,
          but the kernel then lasts as long as its stores take to leave the GPU, and the GEMM queues behind it.

        Two buffer sets alternate, so no cross-GPU barrier is needed: a rank can only start step t+2 (which overwrites
        set t % 2 everywhere) after its GEMM of step t+1, which needed every rank's step-t+1 rows, which every rank sends
        after its own step-t GEMM."""
        from . import _ops

        m, k = x.shape
        # (one batch size per enable_sharded_quant: every block's counter then advances by the same amount per call)
        assert self.sharded_quant_supported(m) and m == self._sq_tokens
        multicast = transport == "multicast"
        assert not multicast or self._sq_mc, "no NVLS multicast mapping on this system"
        b = self._sq_step & 1
        calls = self._sq_step // 2 + 1                       # calls on this buffer set, this one included
        base = b * self._sq_per_set
        off_s, off_c = base + self._sq_codes, base + self._sq_codes + self._sq_scales
        rows = m // self.world
        lo = self.rank * rows
        g = k // 128
        dst = self._sq_mc if multicast else self._sq_buf.data_ptr()
        _ops._quant_bcast(x[lo:lo + rows], dst + base + lo * k, dst + off_s + lo * g * 4, dst + off_c + (lo // 256) * 4,
                          multicast=multicast)
        if not multicast and self.world > 1:
            # after the quant kernel the local counters of the own blocks hold exactly this step's target: they are
            # what is sent behind the data
            spans = [(base + lo * k, rows * k), (off_s + lo * g * 4, rows * g * 4), (off_c + (lo // 256) * 4, (rows // 256) * 4)]
            cur = torch.cuda.current_stream()
            self._sq_ev.record(cur)
            with torch.cuda.stream(self._sq_copy_stream):
                self._sq_copy_stream.wait_event(self._sq_ev)
                for i in range(1, self.world):
                    peer = self._sq_peer[(self.rank - i) % self.world]
                    for off, nbytes in spans:                # codes, scales, then the counters (stream order)
                        peer[off:off + nbytes].copy_(self._sq_buf[off:off + nbytes], non_blocking=True)
        xq = self._sq_buf[base:base + m * k].view(torch.float8_e4m3fn).view(m, k)
        xs = self._sq_buf[off_s:off_s + m * g * 4].view(torch.float32).view(m, g)
        cnt = self._sq_buf[off_c:off_c + (m // 256) * 4].view(torch.int32)
        out = _ops._gemm_gated(xq, self.local.weight_q.t(), xs, self.local.weight_scale.t(), self.local.out_dtype, cnt,
                               calls * _ops.counter_increments(k), self.rank, self.world, whole_panels=not multicast,
                               wait_polls=self.gated_wait_polls)
        self._sq_step += 1
        return out

    def forward(self, x: torch.Tensor, gather_output: bool = True) -> torch.Tensor:
        y = self.forward_local(x)
        return self.gather(y) if gather_output else y

    __call__ = forward


def row_chunk(m: int, world: int) -> int:
    """Rows per rank when [M, K] activations are split for upload: equal chunks (all_gather_into_tensor needs
    them equal), the last ranks' chunks may be partly or wholly beyond M."""
    return -(m // -world) if m > 0 else 0


def gather_row_chunks(buf: torch.Tensor, chunk_rows: int, rank: int, group=None) -> None:
    """In-place all-gather of a row-chunked buffer: `buf` is [world*chunk_rows, ...] contiguous, this rank has
    filled rows [rank*chunk_rows, (rank+1)*chunk_rows); afterwards every rank holds every chunk.  Bytes are
    moved as uint8, so any dtype (e4m3 codes, fp32 scales) goes through NCCL/gloo unchanged."""
    flat = buf.view(torch.uint8).reshape(-1)
    per = flat.numel() // (buf.shape[0] // chunk_rows) if chunk_rows else 0
    if per == 0:
        return
    dist.all_gather_into_tensor(flat, flat[rank * per : (rank + 1) * per], group=group)


class HostStreamedShardedFp8Linear:
    """Host-buffer entry point of the N-sharded layer: x_host [M, K] (pinned, the same batch on every rank) ->
    this rank's [M, N/g] output panel in pinned host memory.

    Uploading the whole batch on every rank would push g copies of x through the host's PCIe root and memory
    (measured at 8 GPUs: 16 ms/step against 11.5 ms on one GPU).  Instead each rank uploads only rows
    [r*ceil(M/g), (r+1)*ceil(M/g)), quantises them with the group-quant op straight into its slice of the full
    [M, K] code / [M, K/128] scale buffers, and one in-place all-gather of codes and one of scales over NVLink
    (1.03 bytes per element instead of 2 over PCIe) completes them on every rank; the local GEMM then produces
    the rank's panel, which leaves over the rank's own PCIe link.  Quantisation is row-local, so the gathered
    codes and scales are bit-identical to quantising the full batch on every rank.  Three streams and two
    buffer sets overlap upload / compute+gather / download of consecutive batches as in HostStreamedFp8Linear."""

    def __init__(self, sharded: ColumnShardedFp8Linear, max_tokens: int):
        layer = sharded.local
        dev = layer.weight_q.device
        self.sh, self.layer = sharded, layer
        self.world, self.rank, self.group = sharded.world, sharded.rank, sharded.group
        k = layer.in_features
        self.max_chunk = row_chunk(max_tokens, self.world)
        rows = self.max_chunk * self.world
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
        self.x_dev = [torch.empty((self.max_chunk, k), dtype=torch.bfloat16, device=dev) for _ in range(2)]
        self.xq = [torch.empty((rows, k), dtype=torch.float8_e4m3fn, device=dev) for _ in range(2)]
        self.xs = [torch.empty((rows, k // 128), dtype=torch.float32, device=dev) for _ in range(2)]
        self.ev_in = [torch.cuda.Event() for _ in range(2)]
        self.ev_cmp = [torch.cuda.Event() for _ in range(2)]
        self.ev_out = [torch.cuda.Event() for _ in range(2)]
        self.count = 0

    def submit(self, x_host: torch.Tensor, y_host: torch.Tensor) -> None:
        """Enqueue one batch (collective: every rank calls it with the same batch, in the same order)."""
        from .gemm import sgl_per_token_group_quant_8bit

        assert x_host.is_pinned() and y_host.is_pinned(), "host buffers must be pinned"
        b = self.count & 1
        m, k = x_host.shape
        chunk = row_chunk(m, self.world)
        assert chunk <= self.max_chunk
        lo = min(self.rank * chunk, m)
        hi = min(lo + chunk, m)
        xq = self.xq[b][: chunk * self.world]
        xs = self.xs[b][: chunk * self.world]
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_cmp[b])          # x_dev[b] no longer read by the quant of batch i-2
            if hi > lo:
                self.x_dev[b][: hi - lo].copy_(x_host[lo:hi], non_blocking=True)
            self.ev_in[b].record(self.s_in)
        with torch.cuda.stream(self.s_cmp):
            self.s_cmp.wait_event(self.ev_in[b])
            self.s_cmp.wait_event(self.ev_out[b])
            if hi > lo:
                sgl_per_token_group_quant_8bit(self.x_dev[b][: hi - lo], xq[lo:hi], xs[lo:hi], 128, 1e-10, -448.0, 448.0,
                                               enable_v2=False)
            if self.world > 1:
                gather_row_chunks(xq, chunk, self.rank, self.group)
                gather_row_chunks(xs, chunk, self.rank, self.group)
            y = self.layer.forward_quantized(xq[:m], xs[:m])
            self.ev_cmp[b].record(self.s_cmp)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_cmp[b])
            y.record_stream(self.s_out)
            y_host.copy_(y, non_blocking=True)
            self.ev_out[b].record(self.s_out)
        self.count += 1

    def synchronize(self) -> None:
        self.s_out.synchronize()
        self.s_cmp.synchronize()
        self.s_in.synchronize()
