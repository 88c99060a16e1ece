"""Row-sharded path on N real GPUs (needs >= 2 devices; the single-GPU driver run skips it).
The host-side N>1 logic is covered on CPU by tests/test_multi_rank_gloo.py."""
import os
import subprocess
import sys

import pytest

import pcsc_eigenvalue_solver_project_b200 as es

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["peer", "nccl"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_row_sharded_power_iteration(world, exchange):
    if es.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    port = 29500 + world + (20 if exchange == "nccl" else 0)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "_mgpu_worker.py")]
    env = dict(os.environ, EIGSOL_B200_EXCHANGE=exchange, EIGSOL_B200_EXPECT_EXCHANGE=exchange,
               EIGSOL_B200_PEER_TIMEOUT_MS="1500")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-4000:]
