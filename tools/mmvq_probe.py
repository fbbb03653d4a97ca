"""Debug/tuning helper (GPU box): time the n=1 MUL_MAT of one shape over rotating weight copies (> L2),
CUDA-graph replayed, CUDA events.  usage: mmvq_probe.py TYPE M K [copies] [reps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads

t, m, k = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
copies = int(sys.argv[4]) if len(sys.argv) > 4 else 0
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 20
kern = b200.Kernels()
wb = m * (k // workloads.BLOCK_ELEMS[t]) * workloads.BLOCK_BYTES[t]
if copies <= 0:
    copies = max(4, int(600e6 // wb) + 1)
ws = [bench.random_blocks_device(torch, kern.device, t, m, k, i) for i in range(copies)]
x = torch.rand((1, k), device=kern.device) * 2 - 1
y = torch.empty((copies, m), dtype=torch.float32, device=kern.device)
nodes = (b200.MulMatNode * copies)(*[b200.mul_mat_node(t, w.data_ptr(), m, k, x.data_ptr(), 1, y[i].data_ptr(), const_weights=True)
                                     for i, w in enumerate(ws)])
kern.mul_mat_batch(nodes); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    kern.mul_mat_batch(nodes)
g.replay(); torch.cuda.synchronize()
sec = bench.timed_loop(torch, g.replay, reps) / (reps * copies)
print(f"{t} m={m} k={k} copies={copies}: {sec*1e6:.2f} us/launch, {wb/sec/1e9:.0f} GB/s on weight bytes "
      f"({wb/sec/1e9/65.514:.1f}% of measured 6551)")
