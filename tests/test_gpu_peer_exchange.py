"""Window close over peer memory (dist.PeerWindowExchange: the export kernel stores into the owners' buffers
over NVLink) against the NCCL all-to-all exchange, on 2 ranks.  Needs two GPUs on one node: skipped on a
single-GPU box."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("hist", [0, 1])
def test_peer_exchange_equals_nccl_exchange_on_two_gpus(hist):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29547", os.path.join(ROOT, "tools", "probe", "peer_exchange_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=dict(os.environ, PEER_HIST=str(hist)))
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    lines = [json.loads(l) for l in out.stdout.splitlines() if l.startswith("{")]
    assert sorted(l["rank"] for l in lines) == [0, 1] and all(l["rows"] > 0 for l in lines)
