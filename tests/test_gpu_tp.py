"""-m gpu, needs two GPUs on one node (skipped otherwise): tensor parallel 2 through torchrun.
scripts/tp_check.py builds one engine per rank, connects the peers' inboxes (CUDA IPC) and checks
that (a) the small geometry reproduces the CPU oracle's greedy tokens with the persistent kernel
running on both ranks and the all-reduce / token exchange inside it, (b) both ranks emit the same
tokens on the 7B geometry."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("nccl_only", ["0", "1"])
def test_tensor_parallel_two_ranks(nccl_only):
    env = dict(os.environ, B200LLM_TP_NCCL_ONLY=nccl_only)
    port = 29600 + int(nccl_only)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", "tp_check.py"), "small"]
    out = subprocess.run(cmd, env=env, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "TP CHECK OK" in out.stdout and "oracle match=True" in out.stdout
