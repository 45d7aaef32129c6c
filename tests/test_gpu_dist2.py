"""GPU, >= 2 devices: the NCCL / NVLink paths of dist.py under torchrun, two ranks spawned by the test itself
(tools/dist_check.py): layer-sharded quantize_targ_layer (codes read from the peer's symmetric-memory buffer or
all-gathered) bit-identical to the single-GPU pass, one tensor in slices, and the batch-sharded eval in sync_stats
mode equal to the un-sharded forward bit for bit.  Skipped on a single-GPU box."""
import os
import socket
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("p2p", ["1", "0"])
def test_two_rank_nccl_paths(p2p):
    env = dict(os.environ, DFQ_DIST_P2P=p2p)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "dist_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "dist_check ok" in out.stdout, out.stdout[-2000:]
    if p2p == "0":
        assert "path all_gather" in out.stdout
