"""Where the fixed ~45 us of an independent-list launch go (tuning library: per-CTA globaltimer stamps).
Runs the 1/8 row share of the Llama-3-70B FFN stack (48 nodes) and the first layers of Llama-2-7B as one
b200_mul_mat_batch each, and prints, relative to the first CTA's start: CTA start spread, node 0 activations ready,
node 0 first stage landed, per-CTA finish of the last node (min / mean / max), producer done.
usage: GGML_B200_KERNELS_LIB=.../libggml_b200_kernels_tuning.so python tools/ramp_probe.py"""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import numpy as np
import torch
import b200, bench, workloads, row_split

kern = b200.Kernels()
dev = kern.device
assert hasattr(kern.lib, "b200_debug_batch_stamps"), "needs the tuning library"
kern.lib.b200_debug_batch_stamps.argtypes = [C.c_void_p]


def run(name, mats, rows):
    weights = [bench.random_blocks_device(torch, dev, mt.type, r, mt.k, 17 + i) for i, (mt, r) in enumerate(zip(mats, rows))]
    xs = [torch.rand(mt.k, device=dev) * 2 - 1 for mt in mats]
    ys = [torch.empty(r, device=dev) for r in rows]
    nodes = (b200.MulMatNode * len(mats))(*[
        b200.mul_mat_node(mt.type, w.data_ptr(), r, mt.k, x.data_ptr(), 1, y.data_ptr(), const_weights=True)
        for mt, w, r, x, y in zip(mats, weights, rows, xs, ys)])
    ws = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(mats))) + 256, dtype=torch.uint8, device=dev)
    fn = lambda: kern.mul_mat_batch(nodes, ws=ws)
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g = bench.capture(torch, fn)
    for _ in range(5): g.replay()
    e0.record()
    for _ in range(100): g.replay()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 100
    nbytes = sum(int(w.numel()) for w in weights)
    n = min(len(mats), 96)
    dbg = torch.zeros((96, 148, 8), dtype=torch.int64, device=dev)
    kern.lib.b200_debug_batch_stamps(C.c_void_p(dbg.data_ptr()))
    torch.cuda.synchronize()
    fn(); fn()                                       # eager, back to back: the second overwrites the first's stamps
    torch.cuda.synchronize()
    kern.lib.b200_debug_batch_stamps(C.c_void_p(0))
    t = dbg.cpu().numpy().astype(np.float64)
    t0 = t[0, :, 6][t[0, :, 6] > 0].min()
    rel = lambda a: (a[a > 0] - t0) / 1e3
    st = lambda a: f"min {a.min():7.2f} mean {a.mean():7.2f} max {a.max():7.2f}" if a.size else "n/a"
    last = n - 1
    print(f"== {name}: {len(mats)} nodes, {nbytes/1e6:.1f} MB, graph replay {ms*1e3:.1f} us/step = {nbytes/ms/1e6:.0f} GB/s; "
          f"at 6100 GB/s the bytes alone take {nbytes/6100e3:.1f} us")
    print("  CTA start            ", st(rel(t[0, :, 6])))
    print("  node 0 acts ready    ", st(rel(t[0, :, 0])))
    print("  node 0 first stage   ", st(rel(t[0, :, 4])))
    print("  node 0 warp0 done    ", st(rel(t[0, :, 5])))
    print(f"  node {last} warp0 done   ", st(rel(t[last, :, 5])))
    print("  producer issued last ", st(rel(t[0, :, 7])))
    # per node, warp 0 of every CTA: previous node done -> activations seen -> first weight stage seen -> done
    ok = (t[1:n, :, 0] > 0) & (t[0:n - 1, :, 5] > 0) & (t[1:n, :, 4] > 0) & (t[1:n, :, 5] > 0)
    w_acts = np.where(ok, t[1:n, :, 0] - t[0:n - 1, :, 5], np.nan) / 1e3
    w_wts = np.where(ok, t[1:n, :, 4] - t[1:n, :, 0], np.nan) / 1e3
    work = np.where(ok, t[1:n, :, 5] - t[1:n, :, 4], np.nan) / 1e3
    print(f"  per node (mean over CTAs and nodes 1..): wait for activations {np.nanmean(w_acts):.3f} us, then for the first weight stage "
          f"{np.nanmean(w_wts):.3f} us, consume {np.nanmean(work):.3f} us")
    for i in range(1, min(n, 7)):
        print(f"    node {i}: acts wait {np.nanmean(w_acts[i-1]):.3f}  weights wait {np.nanmean(w_wts[i-1]):.3f}  consume {np.nanmean(work[i-1]):.3f}")
    # cycle accumulators (rows 93 .. 95 of the table): fractions of the thread's total
    if n <= 90:
        pr, c0, c9 = t[95], t[94], t[93]
        f = lambda num, den: float(np.mean(num[den > 0] / den[den > 0])) * 100
        print(f"  producer: {f(pr[:, 0], pr[:, 1]):.1f} % of its time waiting for a free stage (ring full)")
        for nm, c in (("consumer warp 0", c0), ("consumer warp 9", c9)):
            print(f"  {nm}: waiting for weights {f(c[:, 0], c[:, 4]):.1f} %, group barrier {f(c[:, 1], c[:, 4]):.1f} %, "
                  f"activations {f(c[:, 2], c[:, 4]):.1f} %, layout barrier {f(c[:, 3], c[:, 4]):.1f} % of {np.mean(c[:, 4]) / 1.965e3:.1f} us")
    fin = rel(t[last, :, 5])
    print(f"  finish spread over CTAs: p5 {np.percentile(fin,5):.2f} p50 {np.percentile(fin,50):.2f} p95 {np.percentile(fin,95):.2f}", flush=True)


m70 = workloads.llama3_70b_ffn(n_layers=16)
run("70B FFN, 1/8 rows", m70, [row_split.split_bounds(mt.m, 8)[1] for mt in m70])
run("70B FFN, 1/2 rows (first 16 nodes)", m70[:48], [row_split.split_bounds(mt.m, 2)[1] for mt in m70[:48]])
m7 = workloads.llama2_7b_stack("q4_K", n_layers=12)
run("llama2-7b, 12 layers", m7, [mt.m for mt in m7])
