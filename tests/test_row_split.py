"""CPU tests of the multi-GPU host logic (row split + all-gather), world size 2 over gloo.
The per-rank compute is played by the oracle here (the GPU kernels are row-independent, which
test_gpu_parity.py checks bit-exactly): what is tested is the partition and the exchange."""
import os
import socket

import numpy as np
import pytest

import row_split
from oracle_lib import Oracle, random_blocks


def test_split_bounds_match_reference_rounding():
    assert row_split.split_bounds(28672, 2) == [0, 14336, 28672]
    assert row_split.split_bounds(28672, 8) == [0, 3584, 7168, 10752, 14336, 17920, 21504, 25088, 28672]
    assert row_split.split_bounds(8192, 4) == [0, 2048, 4096, 6144, 8192]
    # boundaries are rounded DOWN to 128 rows; the last rank takes the rest
    b = row_split.split_bounds(32000, 3)
    assert b[0] == 0 and b[-1] == 32000 and all(x % 128 == 0 for x in b[:-1]) and b == sorted(b)
    assert b[1] == 32000 // 3 - (32000 // 3) % 128
    # uneven tensor_split, a rank with zero weight owns nothing
    b = row_split.split_bounds(4096, 3, [3, 0, 1])
    assert b == [0, 3072, 3072, 4096]
    # small matrices: everything lands on the last rank (rounding), never lost
    assert row_split.split_bounds(100, 4) == [0, 0, 0, 0, 100]
    with pytest.raises(ValueError):
        row_split.split_bounds(10, 2, [0, 0])


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, t, m, k, n, tsplit, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = Oracle()
        w = random_blocks(t, m, k, seed=5)                       # every rank derives the same matrix ...
        x = np.random.default_rng(6).uniform(-1, 1, size=(n, k)).astype(np.float32)
        bounds = row_split.split_bounds(m, world, tsplit)
        lo, hi = bounds[rank], bounds[rank + 1]
        out = torch.zeros((n, m), dtype=torch.float32)
        if hi > lo:                                              # ... and computes only its row block
            out[:, lo:hi] = torch.from_numpy(orc.mul_mat(t, w[lo:hi], hi - lo, k, x, mode=1))
        row_split.gather_rows(dist, out, bounds, rank)
        full = orc.mul_mat(t, w, m, k, x, mode=1)
        q.put((rank, bool(np.array_equal(out.numpy().view(np.uint32), full.view(np.uint32)))))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("t,m,k,n,tsplit", [("q4_K", 512, 1024, 1, None), ("q4_K", 640, 1024, 1, [3, 2]),
                                            ("q6_K", 512, 512, 4, None), ("q4_0", 300, 256, 1, None)])
def test_row_split_all_gather_world2_gloo(t, m, k, n, tsplit):
    """2 ranks: split result == 1-rank result, bit for bit, on every rank (equal, uneven, n > 1, and
    the degenerate split where rounding leaves rank 0 empty)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, t, m, k, n, tsplit, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    got = sorted(q.get(timeout=5) for _ in range(2))
    assert got == [(0, True), (1, True)]
