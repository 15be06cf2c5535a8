"""Multi-GPU correctness under pytest: launches tests/_multigpu_worker.py with torchrun on 2 GPUs (and on 4 / 8 when the
box has them) and requires every fused / sharded form of the N-sharded layer to equal the plain per-rank layer call bit
for bit — fused GEMM->all-gather (unicast and NVLS multicast), row-sharded quantisation + gated GEMM (copy-engine and
multicast transports, consecutive steps, a different rank late each step) and the host-buffer pipeline.  Skipped on a
one-GPU box.  The file sorts near the end on purpose: it spawns processes that take every GPU."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_forms_match_the_per_rank_layer(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "_multigpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, f"rc={r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"
    res = json.loads(lines[-1])
    assert res["failed_on_some_rank"] == 0
    for key in ("nccl_gather_equals_unsharded_layer", "fused_gather_unicast", "sharded_quant_copy", "host_streamed_sharded"):
        assert res[key] is True, res
