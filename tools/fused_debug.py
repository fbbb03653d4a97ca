import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), os.path.join(ROOT, "tests")]
import numpy as np, torch
import b200, oracle_lib
t, m, k = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
kern = b200.Kernels(); orc = oracle_lib.Oracle()
w = oracle_lib.random_blocks(t, m, k, seed=3 * m + k)
x = np.random.default_rng(m + k).uniform(-1, 1, size=(1, k)).astype(np.float32)
want = orc.mul_mat(t, w, m, k, x, mode=1).reshape(-1)
wd, xd = kern.to_device(w.reshape(-1)), kern.to_device(x)
for it in range(3):
    got = kern.mul_mat(t, wd, m, k, xd).cpu().numpy().reshape(-1)
    bad = np.nonzero(np.abs(got - want) > 1e-3 * (1 + np.abs(want)))[0]
    print("iter", it, "bad rows:", len(bad), bad[:40])
