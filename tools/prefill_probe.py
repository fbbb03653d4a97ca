"""GPU box: batched MUL_MAT (prefill) TFLOP/s per format / shape / n for the two tensor-core routes:
bf16 (mmq_tc.cuh, default) and the q8_1 int8 route (mmq_i8.cuh, option mmq_route = 2)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads
kern = b200.Kernels()
shapes = [("q4_0", 11008, 4096), ("q8_0", 11008, 4096), ("q4_K", 4096, 4096), ("q4_K", 14336, 4096), ("q5_K", 4096, 11008), ("q6_K", 4096, 14336)]
for n in (32, 128, 512, 2048):
    for t, m, k in shapes:
        mt = workloads.Mat("x", t, m, k)
        w = bench.random_blocks_device(torch, kern.device, t, m, k, 7)
        x = torch.rand((n, k), device=kern.device) * 2 - 1
        y = torch.empty((n, m), dtype=torch.float32, device=kern.device)
        res = {}
        for route in (1, 2):
            if route == 2 and t == "q6_K":
                continue
            prev = kern.set_option("mmq_route", route)
            fn = lambda: kern.mul_mat(t, w, m, k, x, out=y)
            for _ in range(3):
                fn()
            sec = bench.timed_loop(torch, fn, 5) / 5
            kern.set_option("mmq_route", prev)
            res[route] = mt.flops(n) / sec / 1e12
        print(f"n={n:5d} {t} {m}x{k}: bf16 {res[1]:7.1f} TFLOP/s" + (f"   int8 {res[2]:7.1f} TOP/s" if 2 in res else ""), flush=True)
        del w, x, y
