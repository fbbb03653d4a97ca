"""Debug helper: one batched MUL_MAT through the C ABI, NMSE against the oracle (GPU box only)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), os.path.join(ROOT, "tests")]
import b200, oracle_lib

t = sys.argv[1] if len(sys.argv) > 1 else "q4_0"
m, k, n = (int(v) for v in (sys.argv[2:5] if len(sys.argv) > 4 else (256, 512, 64)))
kern = b200.Kernels(); orc = oracle_lib.Oracle()
w = oracle_lib.random_blocks(t, m, k, seed=1)
x = np.random.default_rng(3).uniform(-1, 1, size=(n, k)).astype(np.float32)
import torch
got = kern.mul_mat_q(t, kern.to_device(w.reshape(-1)), m, k, kern.to_device(x))
torch.cuda.synchronize()
got = got.cpu().numpy().reshape(n, m)
print("nmse", t, m, k, n, orc.nmse(orc.mul_mat(t, w, m, k, x, mode=0), got))
