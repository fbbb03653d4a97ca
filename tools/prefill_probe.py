"""Tuning aid (GPU box): the batched MUL_MAT figures of bench.py (n = 512, Llama-2-7B FFN shapes) + n = 2048."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench
kern = b200.Kernels()
for n in (512, 2048):
    print(n, bench.measure_prefill(torch, kern, n=n), flush=True)
