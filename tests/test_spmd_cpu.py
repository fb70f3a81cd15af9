"""SPMD executor over gloo with 2 and 3 ranks (host backend): multi-process plumbing without GPUs."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 3])
def test_spmd_host_backend_over_gloo(world):
    port = 29500 + world + (os.getpid() % 500)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(HERE, "mgpu", "spmd_worker.py"), "--mode", "cpu"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "checks passed" in p.stdout


@pytest.mark.parametrize("faults", ["0", "1"])
def test_spmd_random_queries_match_the_oracle(faults):
    """The SPMD executor on random pipelines (position-dependent operators, single-partition stages over
    cross-product edges, empty partitions): no rank may wait in a collective the others never enter, and
    the gathered result must equal LocalDebug's."""
    env = dict(os.environ, FUZZ_FAULTS=faults)              # "1": a transient failure on one rank in some stages of every job
    last = None
    for attempt in range(2):                            # a busy rendezvous port is not a property of the executor
        port = 29650 + (os.getpid() % 300) + 311 * attempt + 7 * int(faults)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
               "--master-addr", "127.0.0.1", "--master-port", str(port),
               os.path.join(HERE, "mgpu", "spmd_fuzz_worker.py"), "0", "45" if faults == "0" else "25"]
        last = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        if "MISMATCH" in last.stdout:
            break                                       # a wrong result is never retried
        if last.returncode == 0:
            break
    assert last.returncode == 0, last.stdout[-3000:] + last.stderr[-3000:]
    assert "mismatches 0" in last.stdout and "MISMATCH" not in last.stdout
