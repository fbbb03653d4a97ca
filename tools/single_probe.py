"""GPU box: one MUL_MAT per call (a dependent chain's cost): the per-node fused kernel vs the persistent kernel
with a one-node list (option batch_single = 1), rotating weight copies > L2, CUDA graph of the calls."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads
kern = b200.Kernels()
for t, m, k in [("q4_K", 4096, 4096), ("q4_K", 11008, 4096), ("q4_K", 4096, 11008), ("q6_K", 4096, 4096)]:
    mt = workloads.Mat("x", t, m, k)
    copies = max(4, int(600e6 // mt.weight_bytes) + 1)
    ws = [bench.random_blocks_device(torch, kern.device, t, m, k, i) for i in range(copies)]
    x = torch.rand((copies, k), device=kern.device) * 2 - 1
    y = torch.empty((copies, m), dtype=torch.float32, device=kern.device)
    ones = [(b200.MulMatNode * 1)(b200.mul_mat_node(t, w.data_ptr(), m, k, x[i].data_ptr(), 1, y[i].data_ptr(), const_weights=True)) for i, w in enumerate(ws)]
    wsz = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(ones[0], 1)) + 256, dtype=torch.uint8, device=kern.device)
    for mode in ("fused", "persistent-1"):
        prev = kern.set_option("batch_single", 1 if mode == "persistent-1" else -1)
        def run():
            for o in ones: kern.mul_mat_batch(o, ws=wsz)
        run(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g): run()
        g.replay()
        sec = bench.timed_loop(torch, g.replay, 10) / 10 / copies
        kern.set_option("batch_single", prev)
        print(f"{t} {m}x{k} {mode:13s}: {sec*1e6:6.2f} us per MUL_MAT, {mt.algorithmic_bytes(1)/sec/1e9:7.1f} GB/s", flush=True)
