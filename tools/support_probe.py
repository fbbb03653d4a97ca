"""GPU box: bandwidth of the support kernels (dequantize_row_* -> f32/f16, quantize_row_q8_1, GET_ROWS) against
the measured HBM peak.  Bytes = bytes read + bytes written (both algorithmic)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads

kern = b200.Kernels()
peak, _ = bench.hbm_peak()
out = {}
m, k = 32000, 4096                      # the Llama-2 embedding / output matrix
for t in b200.QUANT_TYPES:
    w = bench.random_blocks_device(torch, kern.device, t, m, k, 3)
    for od, osz in (("f32", 4), ("f16", 2)):
        fn = lambda: kern.dequantize_row(t, w, m * k, od)
        for _ in range(3): fn()
        sec = bench.timed_loop(torch, fn, 10) / 10
        byts = w.numel() + m * k * osz
        out[f"dequantize_row_{t}->{od}"] = {"GBps": round(byts / sec / 1e9, 1), "frac_of_measured_peak": round(byts / sec / 1e9 / peak, 3)}
    rows = torch.randint(0, m, (8192,), dtype=torch.int32, device=kern.device)
    fn = lambda: kern.get_rows(t, w, k, rows)
    for _ in range(3): fn()
    sec = bench.timed_loop(torch, fn, 10) / 10
    byts = 8192 * (b200.row_bytes(t, k) + 4 * k)
    out[f"get_rows_{t}(8192 rows)"] = {"GBps": round(byts / sec / 1e9, 1), "frac_of_measured_peak": round(byts / sec / 1e9 / peak, 3)}
    # the same call is 36 us of GPU work behind a Python wrapper: replayed from a CUDA graph (5 calls per replay) the host is out of it
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(5): keepalive = fn()
    g.replay()
    sec = bench.timed_loop(torch, g.replay, 10) / 50
    out[f"get_rows_{t}(8192 rows, graph)"] = {"GBps": round(byts / sec / 1e9, 1), "frac_of_measured_peak": round(byts / sec / 1e9 / peak, 3)}
    del g
    del w
x = torch.rand((8192, 8192), device=kern.device) * 2 - 1
fn = lambda: kern.quantize_row_q8_1(x)
for _ in range(3): fn()
sec = bench.timed_loop(torch, fn, 10) / 10
byts = x.numel() * 4 + x.numel() // 32 * 36
out["quantize_row_q8_1(8192x8192)"] = {"GBps": round(byts / sec / 1e9, 1), "frac_of_measured_peak": round(byts / sec / 1e9 / peak, 3)}
for kk, v in out.items():
    print(f"{kk:40s} {v['GBps']:8.1f} GB/s  {v['frac_of_measured_peak'] * 100:5.1f}% of measured HBM peak")
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "support_probe.json"), "w"), indent=1)
