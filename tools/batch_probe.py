"""Tuning aid (GPU box): time the persistent list kernel on the Llama-2-7B Q4_K decode stack (bench.py's workload).
usage: batch_probe.py [reps] [wtype]      (reps = 0: launch once and exit, for ncu)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "ggml-gfx906_b200"), ROOT]
import torch
import b200, bench, workloads

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
wtype = sys.argv[2] if len(sys.argv) > 2 else "q4_K"
kern = b200.Kernels()
mats = workloads.llama3_8b_q4_k_m_stack() if wtype == "l3mix" else workloads.llama2_7b_stack(wtype)
stack = bench.DecodeStack(kern, mats)
stack.compute(); torch.cuda.synchronize()
if reps == 0:
    stack.compute(); torch.cuda.synchronize()
    sys.exit(0)
c0 = kern.launch_count()
stack.compute(); torch.cuda.synchronize()
launches = kern.launch_count() - c0
sec = bench.timed_loop(torch, stack.compute, reps) / reps
print(f"{wtype} stack: {launches} launches/step, {sec*1e3:.3f} ms/step, {stack.local_alg_bytes/sec/1e9:.0f} GB/s algorithmic "
      f"({stack.local_alg_bytes/sec/1e9/65.517:.1f}% of measured 6551.7), {1/sec:.0f} tok/s", flush=True)
