"""Fixed cost per launch vs per node of b200_mul_mat_batch on one GPU: the 1/world row share of the Llama-3-70B FFN
stack with 4, 8, 16, 32 layers (12 .. 96 nodes); a linear fit over the node count separates the two.
usage: python tools/fixed_cost_probe.py [reps]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ggml-gfx906_b200")); sys.path.insert(0, ROOT)
import numpy as np
import torch
import b200, workloads, row_split, bench

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
kern = b200.Kernels(); dev = kern.device
for world in (8, 2):
    pts = []
    for layers in (4, 8, 16, 32):
        mats = workloads.llama3_70b_ffn(n_layers=layers)
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
        for _ in range(5): g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps): g.replay()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / reps * 1e3
        nbytes = sum(int(w.numel()) for w in weights)
        pts.append((len(mats), us, nbytes))
        print(f"world={world} nodes={len(mats)} {us:.1f} us/step {nbytes/us/1e3:.0f} GB/s", flush=True)
        del weights, xs, ys, ws, g
        torch.cuda.empty_cache()
    n = np.array([p[0] for p in pts], dtype=np.float64); t = np.array([p[1] for p in pts])
    slope, icpt = np.polyfit(n, t, 1)
    per_node_bytes = pts[0][2] / pts[0][0]
    print(f"world={world}: fixed per launch {icpt:.1f} us, per node {slope:.3f} us for {per_node_bytes/1e6:.2f} MB = {per_node_bytes/slope/1e3:.0f} GB/s marginal", flush=True)
