"""Multi-GPU worker (one process per GPU, launched by torchrun from tests/test_gpu_peers.py or by hand):
b200_mul_mat_batch_peers -- the row-split list with the gather fused into the persistent launch -- against the same list
computed un-split on one GPU.  Every rank must end with every row of every output, bit for bit (each row is computed
by exactly one rank with the same kernel and K order), for an independent list and for a dependent FFN chain
(gate/up -> down -> next layer, the gathered vectors feeding the next node inside the same launch), eager, repeated
(the counters' epochs advance) and replayed from a CUDA graph.  Prints PEERS_OK <rank> on success."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "ggml-gfx906_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch
import torch.distributed as dist

import b200
import row_split
from oracle_lib import random_blocks, scale_blocks, Oracle


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    kern = b200.Kernels(device=local)
    dev = kern.device
    orc = Oracle()
    wtype = os.environ.get("PEERS_TYPE", "q4_K")
    hidden, ffn, layers = 1024, 3072, 3

    # ---- the list: per layer gate, up (ffn x hidden) read h; down (hidden x ffn) reads up; the next layer reads down
    specs, cur = [], "h0"
    for l in range(layers):
        specs += [(ffn, hidden, cur, f"g{l}"), (ffn, hidden, cur, f"u{l}"), (hidden, ffn, f"u{l}", f"d{l}")]
        cur = f"d{l}"
    h0 = np.random.default_rng(3).uniform(-1, 1, size=(1, hidden)).astype(np.float32)
    vals, weights = {"h0": h0}, []
    for i, (m, k, src, dst) in enumerate(specs):                      # calibrated to unit gain (see test_gpu_parity._decoder_chain)
        w = random_blocks(wtype, m, k, seed=900 + i)
        y = orc.mul_mat(wtype, w, m, k, vals[src], mode=1)
        scale_blocks(wtype, w, 1.0 / max(float(np.sqrt(np.mean(y.astype(np.float64) ** 2))), 1e-30))
        vals[dst] = orc.mul_mat(wtype, w, m, k, vals[src], mode=1)
        weights.append(w)

    # ---- symmetric region: [ctrl | h0 | every output vector, whole], identical layout on every rank
    off, pos = {}, 0
    for name, n in [("h0", hidden)] + [(dst, m) for (m, k, src, dst) in specs]:
        off[name] = pos
        pos += (4 * n + 127) // 128 * 128
    ctrl = int(kern.lib.b200_peer_ctrl_bytes())
    region = row_split.SymmetricRegion(dist, pos, dev, ctrl)
    pg = region.peer_group()
    base = region.bases[rank] + ctrl

    def vec(name, n):
        return region.data[off[name]: off[name] + 4 * n].view(torch.float32)

    vec("h0", hidden).copy_(torch.from_numpy(h0.reshape(-1)))
    bounds = [row_split.split_bounds(m, world) for (m, k, src, dst) in specs]
    wdev = [kern.to_device(np.ascontiguousarray(w[b[rank]:b[rank + 1]]).reshape(-1)) for w, b in zip(weights, bounds)]

    # ---- reference: the un-split list on this GPU alone (chain: one launch, dependencies inside)
    full_w = [kern.to_device(w.reshape(-1)) for w in weights]
    ref_bufs = {"h0": kern.to_device(h0)}
    for (m, k, src, dst) in specs:
        ref_bufs[dst] = torch.zeros((1, m), dtype=torch.float32, device=dev)
    ref_nodes = (b200.MulMatNode * len(specs))(*[
        b200.mul_mat_node(wtype, w.data_ptr(), m, k, ref_bufs[src].data_ptr(), 1, ref_bufs[dst].data_ptr())
        for (m, k, src, dst), w in zip(specs, full_w)])
    kern.mul_mat_batch(ref_nodes)
    torch.cuda.synchronize()
    ref = {dst: ref_bufs[dst].cpu().numpy().reshape(-1).copy() for (_, _, _, dst) in specs}

    def make_nodes(chain):
        xs = {}
        nodes = []
        for i, ((m, k, src, dst), b, w) in enumerate(zip(specs, bounds, wdev)):
            if chain:
                xptr = base + off[src]
            else:                                   # independent list: every node reads a private copy of the reference run's input
                xs[i] = ref_bufs[src].clone()
                xptr = xs[i].data_ptr()
            nodes.append(b200.mul_mat_node(wtype, w.data_ptr(), b[rank + 1] - b[rank], k, xptr, 1,
                                           base + off[dst] + 4 * b[rank], const_weights=True))
        return (b200.MulMatNode * len(nodes))(*nodes), xs

    def check(tag):
        torch.cuda.synchronize()
        dist.barrier()
        for (m, k, src, dst) in specs:
            got = vec(dst, m).cpu().numpy()
            if not np.array_equal(got.view(np.uint32), ref[dst].view(np.uint32)):
                bad = np.nonzero(got.view(np.uint32) != ref[dst].view(np.uint32))[0]
                raise AssertionError(f"rank {rank} {tag}: {dst} differs at {len(bad)} of {m} rows, first {bad[:5]}")
        dist.barrier()

    def clear():
        torch.cuda.synchronize()
        dist.barrier()
        for (m, k, src, dst) in specs:
            vec(dst, m).zero_()
        torch.cuda.synchronize()
        dist.barrier()

    def say(msg):
        if os.environ.get("PEERS_VERBOSE"):
            print(f"[rank {rank}] {msg}", flush=True)

    # the mapping itself: write a word into every other rank's copy with a plain torch op, read our own back
    probe = region.data[pos - 128: pos - 124].view(torch.int32) if pos >= 128 else None
    ws = None
    for chain in (False, True):
        say(f"chain={chain}")
        nodes, keep = make_nodes(chain)
        ws = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(specs))) + 256, dtype=torch.uint8, device=dev)
        for rep in range(3):                        # eager, the epochs advance
            clear()
            kern.mul_mat_batch_peers(nodes, pg, ws=ws)
            say(f"launched rep {rep}")
            check(f"chain={chain} eager rep {rep}")
            say(f"checked rep {rep}")
        clear()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            kern.mul_mat_batch_peers(nodes, pg, ws=ws)
        for rep in range(4):
            clear()
            g.replay()
            check(f"chain={chain} graph rep {rep}")
        del g
    print(f"PEERS_OK {rank}", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)                                     # (see bench.py: tearing NCCL down after graph use can hang)


if __name__ == "__main__":
    main()
