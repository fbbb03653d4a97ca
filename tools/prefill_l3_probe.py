"""GPU box: BASELINE configs[3] prefill -- one Llama-3-8B layer of the Q4_K_M-like mix (+ the output matrix) as batched
MUL_MATs with n = 2048 (and 512) activation columns: TFLOP/s per matrix and for the layer as one b200_mul_mat_batch list."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads
kern = b200.Kernels()
mats = workloads.llama3_8b_q4_k_m_stack(n_layers=1)
out = {}
for n in (512, 2048):
    ws_, xs, ys = [], [], []
    for i, mt in enumerate(mats):
        w = bench.random_blocks_device(torch, kern.device, mt.type, mt.m, mt.k, 70 + i)
        x = torch.rand((n, mt.k), device=kern.device) * 2 - 1
        y = torch.empty((n, mt.m), dtype=torch.float32, device=kern.device)
        fn = lambda: kern.mul_mat(mt.type, w, mt.m, mt.k, x, out=y)
        for _ in range(2): fn()
        sec = bench.timed_loop(torch, fn, 5) / 5
        out[f"n={n} {mt.name} {mt.type} {mt.m}x{mt.k}"] = round(mt.flops(n) / sec / 1e12, 1)
        print(f"n={n:5d} {mt.name:10s} {mt.type} {mt.m:6d}x{mt.k:<6d}: {mt.flops(n) / sec / 1e12:7.1f} TFLOP/s  {sec * 1e6:8.1f} us", flush=True)
        ws_.append(w); xs.append(x); ys.append(y)
    layer = [i for i, mt in enumerate(mats) if mt.name != "output"]
    nodes = (b200.MulMatNode * len(layer))(*[b200.mul_mat_node(mats[i].type, ws_[i].data_ptr(), mats[i].m, mats[i].k, xs[i].data_ptr(), n, ys[i].data_ptr(), const_weights=True) for i in layer])
    wsz = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(layer))) + 256, dtype=torch.uint8, device=kern.device)
    fn = lambda: kern.mul_mat_batch(nodes, ws=wsz)
    for _ in range(2): fn()
    sec = bench.timed_loop(torch, fn, 5) / 5
    fl = sum(mats[i].flops(n) for i in layer)
    out[f"n={n} layer"] = {"ms": round(sec * 1e3, 3), "TFLOPs": round(fl / sec / 1e12, 1), "tok_s_32_layers": round(n / (32 * sec), 0)}
    print(f"n={n:5d} one layer (7 MUL_MATs, list call): {sec * 1e3:.3f} ms = {fl / sec / 1e12:.1f} TFLOP/s -> {n / (32 * sec):.0f} prompt tok/s for 32 layers (GEMMs only)", flush=True)
    del ws_, xs, ys
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "prefill_l3_probe.json"), "w"), indent=1)
