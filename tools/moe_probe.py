"""GPU box: MUL_MAT_ID decode (mixture of experts): n_as experts of m x k, n_used per token, n_tok tokens; GB/s on the expert
bytes actually read.  usage: moe_probe.py [type] [m] [k] [n_as] [n_used] [n_tok]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench

t = sys.argv[1] if len(sys.argv) > 1 else "q4_K"
m, k, n_as, n_used, n_tok = [int(a) for a in sys.argv[2:7]] if len(sys.argv) > 6 else (14336, 4096, 8, 2, 1)
kern = b200.Kernels()
lib = kern.lib
peak, _ = bench.hbm_peak()
rb = b200.row_bytes(t, k)
layers = max(1, int(1.2e9 // (rb * m * n_as)))            # distinct expert stacks: > L2 in total
ws = [bench.random_blocks_device(torch, kern.device, t, m * n_as, k, i) for i in range(layers)]
x = torch.rand((n_tok, 1, k), device=kern.device) * 2 - 1
ids = torch.stack([torch.randperm(n_as, device=kern.device)[:n_used] for _ in range(n_tok)]).to(torch.int32).contiguous()
y = torch.empty((n_tok, n_used, m), dtype=torch.float32, device=kern.device)
A4i, A4s = C.c_int64 * 4, C.c_size_t * 4
ne_a, nb_a = A4i(k, m, n_as, 1), A4s(b200.BLOCK_BYTES[t], rb, rb * m, rb * m * n_as)
ne_b, nb_b = A4i(k, 1, n_tok, 1), A4s(4, 4 * k, 4 * k, 4 * k * n_tok)
ne_i, nb_i = (C.c_int64 * 2)(n_used, n_tok), (C.c_size_t * 2)(4, 4 * n_used)
lib.b200_mul_mat_id_workspace_size.restype = C.c_size_t
wsz = lib.b200_mul_mat_id_workspace_size(b200.TYPE_ID[t], ne_a, ne_b)
wsb = torch.empty(int(wsz) + 256, dtype=torch.uint8, device=kern.device)
lib.b200_mul_mat_id.restype = C.c_int
def step():
    for w in ws:
        rc = lib.b200_mul_mat_id(b200.TYPE_ID[t], C.c_void_p(w.data_ptr()), ne_a, nb_a, C.c_void_p(x.data_ptr()), ne_b, nb_b,
                                 C.c_void_p(ids.data_ptr()), ne_i, nb_i, C.c_void_p(y.data_ptr()), C.c_void_p(wsb.data_ptr()),
                                 C.c_size_t(wsb.numel()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, rc
step(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g): step()
g.replay()
sec = bench.timed_loop(torch, g.replay, 10) / 10
nbytes = layers * min(n_used * n_tok, n_as) * m * rb if n_tok == 1 else layers * n_used * n_tok * m * rb
print(f"MUL_MAT_ID {t} {m}x{k} experts={n_as} used={n_used} tokens={n_tok}: {sec / layers * 1e6:.1f} us per call, "
      f"{nbytes / sec / 1e9:.0f} GB/s on the expert rows read ({nbytes / sec / 1e9 / peak * 100:.1f}% of measured peak)")
