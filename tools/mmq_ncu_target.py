"""GPU box, ncu target: the prefill kernel on the configs[3] shapes at n = 2048 -- Q4_K 14336x4096 (gate / up), Q6_K 4096x14336 (down) --
one call each after a warm-up call (ncu: -k regex:k_mmq_tc -s 2 -c 2)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench

kern = b200.Kernels()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
cases = [("q4_K", 14336, 4096), ("q6_K", 4096, 14336)]
ws = [bench.random_blocks_device(torch, kern.device, t, m, k, i) for i, (t, m, k) in enumerate(cases)]
xs = [torch.rand((n, k), device=kern.device) * 2 - 1 for (_, _, k) in cases]
for rep in range(2):                                   # launches 0, 1: warm-up; 2, 3: the captured ones
    for (t, m, k), w, x in zip(cases, ws, xs):
        y = kern.mul_mat_q(t, w, m, k, x)
torch.cuda.synchronize()
print("done", [tuple(c) for c in cases], n, float(y.abs().mean()))
