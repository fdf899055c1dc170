"""a-6 on real GPUs (needs >= 2 devices; on a single-GPU box the test skips and `bench.py --gpus N` carries the
driver-visible `dp_check`): data-parallel steps against the oracle on the concatenated batch for both exchanges
(fused peer-memory path over CUDA IPC, NCCL all-reduce), ragged shards, per-rank init seeds, replicas bit-identical."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_peer_exchange_matches_nccl_two_ranks():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "dp_check.py"), "--quick"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=280)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert out.stdout.count("-> OK") == 3 and "MISMATCH" not in out.stdout
    assert '"ok": true' in out.stdout
