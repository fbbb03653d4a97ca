"""GPU box: the decode chain (bench.measure_decode_chain) on the Llama-2-7B Q4_K stack, chained vs cut, a few repeats.
With the tuning library (GGML_B200_KERNELS_LIB=.../libggml_b200_kernels_tuning.so) GGML_B200_CHAIN_DBG=1 / 2 time
'wait only' / 'quantize in the launch only' (wrong results, timing experiments)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads
kern = b200.Kernels()
layers = int(os.environ.get("CHAIN_LAYERS", "32"))
mats = workloads.llama2_7b_stack("q4_K", n_layers=layers)
stack = bench.DecodeStack(kern, mats)
peak, _ = bench.hbm_peak()
for _ in range(3):
    stack.compute()
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    stack.compute()
g.replay()
sec = bench.timed_loop(torch, g.replay, 50) / 50
print(f"independent list: {sec*1e3:.4f} ms/pass, {stack.local_alg_bytes/sec/1e9:.1f} GB/s", flush=True)
out = bench.measure_decode_chain(torch, kern, stack, peak, reps=50)
print(json.dumps(out), flush=True)
if hasattr(kern.lib, "b200_debug_batch_stamps"):
    # tuning library: per-CTA globaltimer stamps of the first 96 nodes of one chained pass
    import ctypes as C, numpy as np
    dbg = torch.zeros((96, 148, 8), dtype=torch.int64, device=kern.device)
    kern.lib.b200_debug_batch_stamps.argtypes = [C.c_void_p]
    kern.lib.b200_debug_batch_stamps(C.c_void_p(dbg.data_ptr()))
    out = bench.measure_decode_chain(torch, kern, stack, peak, reps=3)
    kern.lib.b200_debug_batch_stamps(C.c_void_p(0))
    torch.cuda.synchronize()
    t = dbg.cpu().numpy().astype(np.float64)
    t0 = t[t > 0].min()
    t = np.where(t > 0, t - t0, np.nan) / 1e3          # us
    print("mean over CTAs (us): node | flag seen | released | quantized/acts ready | first stage landed | warp0 done | CTA done mean / max")
    m = lambda a: float("nan") if np.isnan(a).all() else float(np.nanmean(a))
    for i in range(0, min(40, len(mats))):
        print(f"{i:3d} | {m(t[i,:,2]):8.2f} | {m(t[i,:,0]):8.2f} | {m(t[i,:,3]):8.2f} | {m(t[i,:,4]):8.2f} | {m(t[i,:,5]):8.2f} | {m(t[i,:,1]):8.2f} / {np.nanmax(t[i,:,1]) if not np.isnan(t[i,:,1]).all() else float('nan'):8.2f}", flush=True)
