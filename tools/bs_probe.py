"""GPU box: BASELINE configs[3] decode at bs = 1 / 4 / 8: the Llama-3-8B Q4_K_M per-layer GEMM graph (all layers,
distinct weights, 4.6 GB >> L2) as ONE b200_mul_mat_batch list per pass; weight-bytes GB/s and tok/s."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads

kern = b200.Kernels()
peak, _ = bench.hbm_peak()
stack_name = os.environ.get("BS_STACK", "l3mix")
mats = workloads.llama3_8b_q4_k_m_stack() if stack_name == "l3mix" else workloads.llama2_7b_stack(stack_name)
ws = [bench.random_blocks_device(torch, kern.device, mt.type, mt.m, mt.k, i) for i, mt in enumerate(mats)]
wbytes = sum(mt.weight_bytes for mt in mats)
out = {}
for n in [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8]:
    xs = [torch.rand((n, mt.k), device=kern.device) * 2 - 1 for mt in mats]
    ys = [torch.empty((n, mt.m), dtype=torch.float32, device=kern.device) for mt in mats]
    nodes = (b200.MulMatNode * len(mats))(*[b200.mul_mat_node(mt.type, w.data_ptr(), mt.m, mt.k, x.data_ptr(), n, y.data_ptr(), const_weights=True)
                                            for mt, w, x, y in zip(mats, ws, xs, ys)])
    wsz = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(mats))) + 256, dtype=torch.uint8, device=kern.device)
    fn = lambda: kern.mul_mat_batch(nodes, ws=wsz)
    fn(); torch.cuda.synchronize()
    c0 = kern.launch_count(); fn(); launches = kern.launch_count() - c0
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): fn()
    g.replay()
    sec = bench.timed_loop(torch, g.replay, 10) / 10
    out[f"bs={n}"] = {"launches": launches, "ms_per_pass": round(sec * 1e3, 3), "weight_GBps": round(wbytes / sec / 1e9, 1),
                      "frac_of_measured_peak": round(wbytes / sec / 1e9 / peak, 3), "tok_per_s": round(n / sec, 1)}
    print(f"{stack_name} bs={n}: {launches} launches/pass, {sec*1e3:.3f} ms/pass, {wbytes/sec/1e9:.0f} GB/s on weight bytes "
          f"({wbytes/sec/1e9/peak*100:.1f}% of measured peak), {n/sec:.0f} tok/s", flush=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "bs_probe.json"), "w"), indent=1)
