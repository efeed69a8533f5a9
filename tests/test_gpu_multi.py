"""Cell-sharded multi-GPU parity: N ranks over NCCL against the oracle on the whole matrix.
Needs >= 2 GPUs (gpurun --gpus 2); skipped on a one-GPU box."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("p2p", ["1", "0"], ids=["peer_memory", "nccl"])
@pytest.mark.parametrize("case", ["multi", "multi_big"])
def test_sharded_matches_oracle(case, p2p):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "multi_worker.py"), case]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env={**os.environ, "CLX_P2P": p2p})
    assert r.returncode == 0 and "MULTI_OK" in r.stdout and "MULTI_SOUP_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
