"""GPU box: a few launches of the batched MUL_MAT on one shape (for ncu).  usage: mmq_once.py TYPE M K N"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench
t, m, k, n = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
kern = b200.Kernels()
w = bench.random_blocks_device(torch, kern.device, t, m, k, 77)
x = torch.rand((n, k), device=kern.device) * 2 - 1
y = torch.empty((n, m), dtype=torch.float32, device=kern.device)
for _ in range(3):
    kern.mul_mat(t, w, m, k, x, out=y)
torch.cuda.synchronize()
