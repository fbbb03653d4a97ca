"""Tuning aid (GPU box): per-CTA globaltimer stamps of consecutive fused decode launches.
usage: fused_timeline.py TYPE M K [launches]"""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch, numpy as np
import b200, bench
t, m, k = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
L = int(sys.argv[4]) if len(sys.argv) > 4 else 6
kern = b200.Kernels()
ws = [bench.random_blocks_device(torch, kern.device, t, m, k, i) for i in range(L)]
x = torch.rand((1, k), device=kern.device) * 2 - 1
y = torch.empty((L, m), dtype=torch.float32, device=kern.device)
dbgs = [torch.zeros((320, 8), dtype=torch.int64, device=kern.device) for _ in range(L)]
def run():
    for i in range(L):
        kern.lib.b200_debug_timestamps(ctypes.c_void_p(dbgs[i].data_ptr()))
        kern.mul_mat(t, ws[i], m, k, x, out=y[i:i+1], const_weights=True)
    kern.lib.b200_debug_timestamps(None)
run(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    run()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
full_dump = [x_.cpu().numpy().reshape(-1) for x_ in dbgs]
d = [x_.cpu().numpy()[:160] for x_ in dbgs]
t0 = min(a[a[:, 0] > 0, 0].min() for a in d)
names = ["cta_start", "dep_ok", "quant_done", "first_slot", "last_slot", "done"]
print("launch  " + "  ".join(f"{n:>22}" for n in names) + "   (us since first CTA start: min / median / max over CTAs)")
for i, a in enumerate(d):
    a = a[a[:, 0] > 0]
    cells = []
    for j in range(6):
        v = (a[:, j] - t0) / 1e3
        cells.append(f"{v.min():6.2f}/{np.median(v):6.2f}/{v.max():6.2f}")
    print(f"{i:6d}  " + "  ".join(f"{c:>22}" for c in cells) + f"   ctas={len(a)}")

a = full_dump[min(2, L - 1)]
iss, ful, arr, obs = a[2048:2048+120], a[2048+128:2048+248], a[2048+256:2048+376], a[2048+384:2048+504]
nz = iss > 0
base = iss[nz].min()
print("CTA 0 of launch 2: slot b: issue / landed (observer) / full seen by warp 0 / last warp arrived (us)")
for b in range(int(nz.sum())):
    print(f"  b={b:3d}  {(iss[b]-base)/1e3:7.2f}  {(obs[b]-base)/1e3:7.2f}  {(ful[b]-base)/1e3:7.2f}  {(arr[b]-base)/1e3:7.2f}")
