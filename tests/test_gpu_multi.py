"""Multi-GPU correctness inside the driver's `pytest -m gpu` run: launches the SPMD worker under
torchrun on every visible GPU (2 at least) and requires its end-of-run marker.  Covers the fused
shuffle + protocol checker, the split key/payload sort, both aggregate merges, NVLS, broadcast
join, k-means all-reduce (both forms), the streamed channels and the public API at N > 1."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:  # noqa: BLE001
        return 0


@pytest.mark.gpu
@pytest.mark.parametrize("nproc", [2, 8])
def test_spmd_worker_gpu(nproc):
    have = _ngpus()
    if have < nproc:
        pytest.skip(f"needs {nproc} GPUs, box has {have}")
    port = 29650 + nproc
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "mgpu", "spmd_worker.py"), "--mode", "gpu"]
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", DRYAD_SHUFFLE_VERIFY="1")
    p = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    tail = (p.stdout + p.stderr)[-4000:]
    assert p.returncode == 0, tail
    assert f"SPMD gpu checks passed on {nproc} ranks" in p.stdout, tail
