"""The fused row-split gather (b200_mul_mat_batch_peers) on real GPUs: needs >= 2 devices, one process per GPU under
torchrun (tests/peers_worker.py does the work and the bit-exact checks)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("wtype", ["q4_K", "q6_K", "q8_0"])
def test_fused_row_split_gather_two_ranks(wtype):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 2
    env = dict(os.environ, PEERS_TYPE=wtype)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + os.getpid() % 400), os.path.join(ROOT, "tests", "peers_worker.py")]
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.count("PEERS_OK") == world, r.stdout[-3000:]
