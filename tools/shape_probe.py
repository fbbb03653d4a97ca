"""GPU box: per-shape decode mat-vec bandwidth (BASELINE configs[2] shapes, plus the Llama-3-70B FFN shapes of
configs[4]): each (format, m, k) as a list of distinct weight copies totalling > 600 MB (>> L2), one
b200_mul_mat_batch call per pass, CUDA events.  GB/s on algorithmic bytes, % of the measured HBM copy peak."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads

kern = b200.Kernels()
peak, _ = bench.hbm_peak()
shapes = [(4096, 4096), (11008, 4096), (4096, 11008), (32000, 4096), (28672, 8192), (8192, 28672)]
types = sys.argv[1:] or ["q4_K", "q5_K", "q6_K"]
out = {}
for t in types:
    for m, k in shapes:
        mt = workloads.Mat("x", t, m, k)
        copies = max(4, int(600e6 // mt.weight_bytes) + 1)
        ws = [bench.random_blocks_device(torch, kern.device, t, m, k, i) for i in range(copies)]
        x = torch.rand((copies, k), device=kern.device) * 2 - 1
        y = torch.empty((copies, m), dtype=torch.float32, device=kern.device)
        nodes = (b200.MulMatNode * copies)(*[b200.mul_mat_node(t, w.data_ptr(), m, k, x[i].data_ptr(), 1, y[i].data_ptr(), const_weights=True)
                                             for i, w in enumerate(ws)])
        wsz = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, copies)) + 256, dtype=torch.uint8, device=kern.device)
        fn = lambda: kern.mul_mat_batch(nodes, ws=wsz)
        for _ in range(3): fn()
        sec = bench.timed_loop(torch, fn, 10) / 10
        gbps = copies * mt.algorithmic_bytes(1) / sec / 1e9
        out[f"{t} {m}x{k}"] = {"copies": copies, "us_per_matvec": round(sec / copies * 1e6, 2), "GBps": round(gbps, 1), "frac_of_measured_peak": round(gbps / peak, 3)}
        print(f"{t:5s} m={m:6d} k={k:6d} x{copies:3d}: {sec / copies * 1e6:7.2f} us/mat-vec  {gbps:7.1f} GB/s  {gbps / peak * 100:5.1f}% of measured peak", flush=True)
        del ws, x, y, wsz
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "shape_probe.json"), "w"), indent=1)
