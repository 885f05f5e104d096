"""Multi-GPU checks (skipped on a single-GPU box): peer-memory fused all-reduce + update, SyncBatchNorm over NVLink,
replicas bit-identical after training steps.  Runs benchmarks/multigpu_check.py under torchrun."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_memory_engine_two_or_more_gpus():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "benchmarks", "multigpu_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert '"test": "SUMMARY"' in r.stdout and '"ok": true' in r.stdout.splitlines()[-1]


def test_nvls_probe_reports():
    from picd_b200.parallel import native_comm
    assert native_comm.nvls_supported() in (-1, 0, 1)
