"""Compute-only time of ONE rank's share of the row-split Llama-3-70B FFN stack (BASELINE configs[4]) on a single GPU,
for world = 1, 2, 4, 8: what b200_mul_mat_batch needs for the 1/world row blocks when nothing is gathered.  Against
the multi-GPU step time of bench.py --gpus N this separates the kernel's small-problem efficiency from the gather.
usage: python tools/shard_probe.py [reps]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-gfx906_b200")); sys.path.insert(0, ROOT)
import torch
import b200, workloads, row_split
import bench

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
kern = b200.Kernels()
dev = kern.device
mats = workloads.llama3_70b_ffn(n_layers=16)
out = {}
for world in (1, 2, 4, 8):
    rows = [row_split.split_bounds(mt.m, world)[1] for mt in mats]
    weights = [bench.random_blocks_device(torch, dev, mt.type, r, mt.k, 17 + i) for i, (mt, r) in enumerate(zip(mats, rows))]
    xs = [torch.rand(mt.k, device=dev) * 2 - 1 for mt in mats]
    ys = [torch.empty(r, device=dev) for r in rows]
    nodes = (b200.MulMatNode * len(mats))(*[
        b200.mul_mat_node(mt.type, w.data_ptr(), r, mt.k, x.data_ptr(), 1, y.data_ptr(), const_weights=True)
        for mt, w, r, x, y in zip(mats, weights, rows, xs, ys)])
    ws = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(mats))) + 256, dtype=torch.uint8, device=dev)
    fn = lambda: kern.mul_mat_batch(nodes, ws=ws)
    for _ in range(5): fn()
    g = bench.capture(torch, fn)
    step = g.replay if g is not None else fn
    for _ in range(5): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    nbytes = sum(int(w.numel()) for w in weights)
    out[f"world={world}"] = {"ms_per_step": round(ms, 4), "weight_GB": round(nbytes / 1e9, 3), "GBps": round(nbytes / ms / 1e6, 1),
                            "rows_per_cta": [round(r / 148, 1) for r in sorted(set(rows))]}
    print(f"world={world}", json.dumps(out[f"world={world}"]), flush=True)
    del weights, xs, ys, ws, g
    torch.cuda.empty_cache()
