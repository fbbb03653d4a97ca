"""Row-split of a weight matrix over the GPUs of one box and the gather of the output row blocks.

Host-side mirror of the reference's split buffer arithmetic: `tensor_split` fractions -> cumulative
boundaries (src/ggml-cuda/ggml-cuda.cu:1060-1074), row_low/row_high rounded down to `rounding` rows
(get_row_split / get_row_rounding, :765-791; 128 = mmq_y on NVIDIA), the last device taking the rest.
The reference gathers the row blocks to the main device with peer copies (:1441-1759); here every rank
ends with the full output through ONE all-gather (torch.distributed: NCCL over NVLink on GPUs, gloo in
the CPU tests).  Each output element is produced by exactly one rank with the same kernel and the same
K order, so an N-rank result is bit-identical to the 1-rank result.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple


def split_bounds(m: int, world: int, tensor_split: Optional[Sequence[float]] = None, rounding: int = 128) -> List[int]:
    """Row boundaries [b_0 = 0, b_1, ..., b_world = m]: rank r owns rows [b_r, b_{r+1})."""
    if world < 1:
        raise ValueError("world must be >= 1")
    if tensor_split is None:
        tensor_split = [1.0] * world
    if len(tensor_split) != world or any(f < 0 for f in tensor_split) or sum(tensor_split) <= 0:
        raise ValueError("tensor_split needs one non-negative weight per rank, not all zero")
    total = float(sum(tensor_split))
    bounds, acc = [], 0.0
    for r in range(world):
        low = int(m * (acc / total))
        bounds.append(low - low % rounding)
        acc += tensor_split[r]
    bounds.append(m)
    bounds[0] = 0
    return bounds


def rank_rows(m: int, world: int, rank: int, tensor_split: Optional[Sequence[float]] = None, rounding: int = 128) -> Tuple[int, int]:
    b = split_bounds(m, world, tensor_split, rounding)
    return b[rank], b[rank + 1]


def gather_rows(dist, out, bounds: Sequence[int], rank: int):
    """All ranks hold their row block in out[..., b_r:b_{r+1}] (out: [n, m], any device); after the call
    every rank holds all of out.  Equal blocks and n == 1: one all_gather_into_tensor in place; otherwise
    blocks are padded to the largest one (still a single collective)."""
    import torch
    world = len(bounds) - 1
    if world == 1:
        return out
    sizes = [bounds[r + 1] - bounds[r] for r in range(world)]
    n = out.shape[0]
    if n == 1 and len(set(sizes)) == 1 and out.is_contiguous():
        flat = out.view(-1)
        dist.all_gather_into_tensor(flat, flat[bounds[rank]:bounds[rank + 1]])
        return out
    big = max(sizes)
    mine = torch.zeros((n, big), dtype=out.dtype, device=out.device)
    mine[:, :sizes[rank]] = out[:, bounds[rank]:bounds[rank + 1]]
    parts = torch.empty((world, n, big), dtype=out.dtype, device=out.device)
    dist.all_gather_into_tensor(parts.view(-1), mine.view(-1))
    for r in range(world):
        if r != rank:
            out[:, bounds[r]:bounds[r + 1]] = parts[r, :, :sizes[r]]
    return out


class SymmetricRegion:
    """One buffer per rank, every rank's copy mapped into every process (NVLink peer mappings), for
    b200_mul_mat_batch_peers: the row blocks a rank computes are stored straight into all copies from inside the
    persistent launch.  Plumbing only: torch.distributed's symmetric memory does the allocation and the handle
    exchange (CUDA VMM / IPC); `bases[r]` is the address of rank r's copy in THIS process.  The first
    `ctrl_bytes` bytes belong to the kernel library (zeroed here); `data` is the rest as a uint8 tensor."""

    def __init__(self, dist, nbytes: int, device, ctrl_bytes: int):
        import torch
        import torch.distributed._symmetric_memory as symm
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.nbytes = (int(nbytes) + ctrl_bytes + 255) // 256 * 256
        self.buf = symm.empty(self.nbytes, dtype=torch.uint8, device=device)
        self.handle = symm.rendezvous(self.buf, dist.group.WORLD.group_name)
        self.bases = [int(p) for p in self.handle.buffer_ptrs]
        assert len(self.bases) == self.world and self.bases[self.rank] == self.buf.data_ptr()
        self.buf.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        self.ctrl_bytes = ctrl_bytes
        self.data = self.buf[ctrl_bytes:]

    def peer_group(self):
        import b200
        return b200.peer_group(self.rank, self.bases, self.nbytes)
