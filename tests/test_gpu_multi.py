"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): launches scripts/ddp_check.py
under torchrun, one rank per GPU over NCCL.  Logs of the same command at 2, 4 and 8 GPUs are committed under
profiles/r02_ddp_check_w{2,4,8}.txt; bench.py runs the core of it (oracle/probe.py) before timing at every world size."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_flat_data_parallel_matches_reference_ddp(world):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29540 + world), os.path.join(ROOT, "scripts", "ddp_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "match the oracle" in out.stdout and "graph replay == eager" in out.stdout
    # per-rank batch 1024 and 128 against the oracle's W-rank restatement and against one GPU on the global batch
    assert out.stdout.count('"ok": true') >= 3 and '"ok": false' not in out.stdout
    assert "a late rank raises on every rank" in out.stdout
