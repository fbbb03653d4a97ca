#!/usr/bin/env python
"""bench.py — the headline measurement of the B200 ggml quantized hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--stack ...]

N = 1 (default workload: BASELINE.json configs[2]): the decode mat-vec (n = 1) over every projection of Llama-2-7B in
Q4_K: 225 matrices, 3.716 GB of quant blocks, far larger than L2, so every step streams its weights from HBM (no L2
flush needed).  A "step" is one token: ONE b200_mul_mat_batch call (one quantize + one persistent launch).
  metric  quant_matvec_GBps = algorithmic bytes of the step / step time
  value   device-resident: weights, activations and outputs in HBM, the step replayed from a CUDA graph, K steps
  e2e     the same step through the ggml plugin boundary, the calls an unmodified ggml program makes:
          ggml_backend_tensor_set_async (pinned host activations, ggml_backend_b200_host_buffer_type) ->
          ggml_backend_graph_compute_async on B200_0 -> ggml_backend_tensor_get_async (pinned host outputs) ->
          ggml_backend_synchronize, every step inside the timed region (weights are the model: uploaded once)
  extras  decode_chain (the same stack with the dependency structure of a real decoder pass, one launch),
          decode_batches (n = 2/4/8), prefill (tensor-core path, with its own roofline against the measured tcgen05
          peaks), support kernels, the CPU baseline (the unmodified reference's ggml-cpu).
N > 1 (default workload: BASELINE configs[4]): 16 layers of the Llama-3-70B FFN (gate, up 28672 x 8192; down 8192 x 28672)
in Q4_K with the ROWS SPLIT over the ranks (strong scaling: the job is the same at every N).  Every rank computes its
row blocks and stores them straight into every rank's copy of the output arena from inside the persistent launch
(b200_mul_mat_batch_peers: the all-gather fused into the kernel, NVLink peer stores + system-scope completion flags);
the per-group NCCL all-gather is kept as the baseline leg.  After the timed regions every rank re-computes a sample of
OTHER ranks' rows from regenerated weights and compares bit for bit.  `--stack llama2-7b` keeps round 1's weak-scaling
run (rows x N).

--impl reference times the reference's own CPU implementation (ggml-cpu from oracle/_ref, all host threads) on a
bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "ggml-gfx906_b200")
if PKG not in sys.path:
    sys.path.insert(0, PKG)

import workloads  # noqa: E402

METRIC = "quant_matvec_GBps"
UNIT = "GB/s"
WORKLOAD = "llama2-7b all projections, Q4_K, decode mat-vec n=1 (BASELINE configs[2])"


def dbg(msg):
    """milestones to stderr when BENCH_DEBUG is set (multi-rank hangs are otherwise silent)"""
    if os.environ.get("BENCH_DEBUG"):
        sys.stderr.write(f"[bench r{os.environ.get('RANK', '0')} {time.perf_counter():.1f}] {msg}\n")
        sys.stderr.flush()


def env_int(name, dflt):
    try:
        return int(os.environ.get(name, dflt))
    except ValueError:
        return dflt


# ------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except ValueError:
                continue
            for nm, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------- synthetic data
F16_FIELDS = {"q4_0": (0,), "q8_0": (0,), "q4_K": (0, 2), "q5_K": (0, 2), "q6_K": (208,)}
SCALE_RANGE = {"q4_0": (0.02, 0.15), "q8_0": (0.001, 0.01), "q4_K": (0.0005, 0.004),
               "q5_K": (0.0003, 0.002), "q6_K": (0.0001, 0.0006)}


def random_blocks_device(torch, dev, wtype, rows, k, seed):
    """Random-quantized weights made on the device: every bit pattern of a block's integer fields is
    legal, so random bytes + sane fp16 scale fields are valid blocks (same recipe as
    tests/oracle_lib.random_blocks, which the parity tests use)."""
    nb = rows * (k // workloads.BLOCK_ELEMS[wtype])
    bs = workloads.BLOCK_BYTES[wtype]
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    raw = torch.randint(0, 256, (nb, bs), dtype=torch.uint8, device=dev, generator=g)
    lo, hi = SCALE_RANGE[wtype]
    for off in F16_FIELDS[wtype]:
        s = (torch.rand(nb, device=dev, generator=g) * (hi - lo) + lo).to(torch.float16)
        if off in (0, 208):
            s = s * (torch.randint(0, 2, (nb,), device=dev, generator=g).to(torch.float16) * 2 - 1)
        raw[:, off:off + 2] = s.view(torch.uint8).view(nb, 2)
    return raw.view(-1)


def scale_blocks_device(torch, wtype, weights, factor):
    """multiply every fp16 scale field of the (flat uint8) block tensor by `factor`, in place"""
    v = weights.view(-1, workloads.BLOCK_BYTES[wtype])
    for off in F16_FIELDS[wtype]:
        f = v[:, off:off + 2].contiguous().view(torch.float16).float() * factor
        v[:, off:off + 2] = f.to(torch.float16).view(torch.uint8).view(-1, 2)


def mul_mat_groups(mats):
    """MUL_MAT groups of a decoder pass = runs of matrices that read the same activations (q/k/v, gate/up) or stand alone
    (wo, down, output): the granularity at which a row-split model has to gather (one gather per group)."""
    groups, i = [], 0
    while i < len(mats):
        nm = mats[i].name.split(".")[-1]
        run = {"wq": 3, "gate": 2}.get(nm, 1)
        groups.append(list(range(i, min(i + run, len(mats)))))
        i += run
    return groups


def chain_plan(mats):
    """(matrix index, source vector, destination vector) with the dependency structure of a decoder pass: the first group
    of a layer reads the layer input, every later group reads the LAST output of the group before it (wo the attention
    stand-in = the last of q/k/v, gate/up read wo, down reads up), the next layer reads down, the output matrix the last down."""
    plan, cur = [], "h0"
    for g in mul_mat_groups(mats):
        for i in g:
            plan.append((i, cur, f"y{i}"))
        cur = f"y{g[-1]}"
    return plan


class DecodeStack:
    """All matrices of the workload resident on one device, activations and the output arena.

    Output arena ("full-vector layout"): matrix i owns m_i floats at off[i]; with the rows split over `world` ranks
    (row_split.split_bounds) rank r's row block sits at off[i] + bounds[i][r] -- on EVERY rank once gathered.  world > 1:
    the arena is a symmetric region (one copy per rank, peer-mapped) for the fused gather of b200_mul_mat_batch_peers;
    the NCCL baseline leg uses a second arena in group-rank-major layout (one all_gather_into_tensor per MUL_MAT group)."""

    def __init__(self, kern, mats, rank=0, world=1, dist=None):
        import torch
        import b200
        import row_split
        self.torch, self.kern, self.mats, self.rank, self.world, self.dist = torch, kern, mats, rank, world, dist
        dev = kern.device
        self.bounds = [row_split.split_bounds(mt.m, world) for mt in mats]
        self.rows = [(b[rank], b[rank + 1]) for b in self.bounds]                   # my row block per matrix
        for b in self.bounds:
            assert len({b[r + 1] - b[r] for r in range(world)}) == 1, "bench shapes split evenly"
        self.weights = [random_blocks_device(torch, dev, mt.type, hi - lo, mt.k, self.weight_seed(i, rank))
                        for i, (mt, (lo, hi)) in enumerate(zip(mats, self.rows))]
        self.groups = mul_mat_groups(mats)
        # ---- activations: one private vector per matrix (the independent list), identical on every rank
        x_off, xo = [], 0
        for mt in mats:
            x_off.append(xo); xo += mt.k
        self.x_off = x_off
        g = torch.Generator(device="cpu"); g.manual_seed(1234)
        self.x_host = (torch.rand(xo, generator=g) * 2 - 1).pin_memory()
        self.x_dev = self.x_host.to(dev)
        # ---- output arena, full-vector layout (+ h0 for the chain at the front)
        self.h0_len = mats[0].k
        off, yo = [], self.h0_len
        for mt in mats:
            off.append(yo); yo += (mt.m + 31) // 32 * 32
        self.off, self.arena_len = off, yo
        self.region = None
        if world > 1:
            ctrl = int(kern.lib.b200_peer_ctrl_bytes())
            try:
                self.region = row_split.SymmetricRegion(dist, 4 * yo, dev, ctrl)
                self.y_dev = self.region.data[:4 * yo].view(torch.float32)
                self.pg = self.region.peer_group()
            except Exception as e:              # noqa: BLE001 - no peer mapping on this box: only the NCCL leg runs
                sys.stderr.write(f"[bench r{rank}] symmetric memory unavailable ({str(e)[:200]}): fused gather leg skipped\n")
                self.region = None
        if self.region is None:
            self.y_dev = torch.zeros(yo, dtype=torch.float32, device=dev)
        gh = torch.Generator(device="cpu"); gh.manual_seed(4321)
        self.y_dev[:self.h0_len].copy_(torch.rand(self.h0_len, generator=gh) * 2 - 1)
        self.y_host = torch.empty(yo, dtype=torch.float32).pin_memory()

        def node(i, xptr, dptr):
            mt, (lo, hi) = mats[i], self.rows[i]
            return b200.mul_mat_node(mt.type, self.weights[i].data_ptr(), hi - lo, mt.k, xptr, 1, dptr, const_weights=True)

        ybase = self.y_dev.data_ptr()
        self.nodes = (b200.MulMatNode * len(mats))(*[
            node(i, self.x_dev.data_ptr() + 4 * x_off[i], ybase + 4 * (off[i] + self.rows[i][0])) for i in range(len(mats))])
        self.ws = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(self.nodes, len(mats))) + 256,
                              dtype=torch.uint8, device=dev)
        # ---- the same matrices with the dependency structure of a decoder pass (inputs = arena vectors)
        self.plan = chain_plan(mats)
        src_off = {"h0": 0, **{f"y{i}": off[i] for i in range(len(mats))}}
        self.chain_nodes = (b200.MulMatNode * len(mats))(*[
            node(i, ybase + 4 * src_off[src], ybase + 4 * (off[i] + self.rows[i][0])) for i, src, _ in self.plan])
        self.chain_calibrated = False
        # ---- NCCL baseline leg (world > 1): group-rank-major arena, one all-gather per MUL_MAT group
        if world > 1:
            self.nccl_segs, self.nccl_nodes, pos = [], [], 0
            gsz = [sum(self.rows[i][1] - self.rows[i][0] for i in g_) for g_ in self.groups]
            self.y_nccl = torch.zeros(world * sum(gsz), dtype=torch.float32, device=dev)
            for g_, sz in zip(self.groups, gsz):
                seg = self.y_nccl[pos: pos + world * sz]
                mine, p2, sub = seg[rank * sz:(rank + 1) * sz], 0, []
                for i in g_:
                    sub.append(node(i, self.x_dev.data_ptr() + 4 * x_off[i], mine.data_ptr() + 4 * p2))
                    p2 += self.rows[i][1] - self.rows[i][0]
                self.nccl_segs.append((seg, mine))
                self.nccl_nodes.append((b200.MulMatNode * len(sub))(*sub))
                pos += world * sz
        self.local_weight_bytes = sum(int(w.numel()) for w in self.weights)
        self.local_alg_bytes = sum((hi - lo) * (mt.k // workloads.BLOCK_ELEMS[mt.type]) * workloads.BLOCK_BYTES[mt.type]
                                   + mt.k * 4 + (hi - lo) * 4 for mt, (lo, hi) in zip(mats, self.rows))
        self.x_bytes, self.y_bytes = 4 * xo, 4 * (yo - self.h0_len)

    @staticmethod
    def weight_seed(i, rank):
        return 0xB200 + 7919 * i + rank

    # ------------------------------------------------------------------ one step
    def compute(self):
        """Every MUL_MAT of the step in one C-ABI call: one quantize + one persistent launch; with the rows split over
        ranks the launch also stores every finished row into every rank's arena (the gather, fused)."""
        if self.world > 1 and self.region is not None:
            self.kern.mul_mat_batch_peers(self.nodes, self.pg, ws=self.ws)
        elif self.world > 1:
            self.compute_nccl()
        else:
            self.kern.mul_mat_batch(self.nodes, ws=self.ws)

    def compute_chain(self):
        if self.world > 1:
            self.kern.mul_mat_batch_peers(self.chain_nodes, self.pg, ws=self.ws)
        else:
            self.kern.mul_mat_batch(self.chain_nodes, ws=self.ws)

    def compute_nccl(self):
        """baseline: one persistent launch + one NCCL all-gather per MUL_MAT group (what a row-split model does without the fusion)"""
        for sub, (seg, mine) in zip(self.nccl_nodes, self.nccl_segs):
            self.kern.mul_mat_batch(sub, ws=self.ws)
            self.dist.all_gather_into_tensor(seg, mine)

    def calibrate_chain(self):
        """Random blocks amplify (|y| ~ 30 |x| per node): without the normalisations of a real model the chain overflows
        after a few layers.  Stand-in for them: every matrix's fp16 scale fields are scaled, once, to unit output RMS
        (row-split: every rank applies the same factor, computed from the gathered vector)."""
        if self.chain_calibrated:
            return
        import b200
        torch, kern = self.torch, self.kern
        for idx, (i, _, _) in enumerate(self.plan):
            mt = self.mats[i]
            one = (b200.MulMatNode * 1)(self.chain_nodes[idx])
            for it in range(2):                       # measure the gain, rescale, recompute (the next node reads the calibrated output)
                if self.world > 1:
                    kern.mul_mat_batch_peers(one, self.pg, ws=self.ws)
                    torch.cuda.synchronize(); self.dist.barrier()
                else:
                    kern.mul_mat_batch(one, ws=self.ws)
                if it == 0:
                    rms = float(self.y_dev[self.off[i]: self.off[i] + mt.m].double().pow(2).mean().sqrt())
                    scale_blocks_device(torch, mt.type, self.weights[i], 1.0 / max(rms, 1e-30))
        self.chain_calibrated = True

    # ------------------------------------------------------------------ C-ABI end to end (extra leg)
    def _e2e_setup(self, chunks):
        """Split the node list into `chunks` consecutive sub-lists with their activation / output ranges, so
        that the host->device copy of sub-list c+1 and the device->host copy of sub-list c-1 overlap the
        persistent launch of sub-list c (three streams, events in between)."""
        import b200
        torch = self.torch
        n = len(self.mats)
        total = sum(int(w.numel()) for w in self.weights)
        edge = float(os.environ.get("BENCH_E2E_EDGE", "0.1"))
        if chunks >= 4 and 0.0 < edge < 0.25:
            cuts = [edge + (1.0 - 2.0 * edge) * j / (chunks - 2) for j in range(chunks - 1)]
        else:
            cuts = [j / chunks for j in range(1, chunks)]
        edges, acc, nxt = [0], 0, 0
        for i, w in enumerate(self.weights):
            acc += int(w.numel())
            while nxt < len(cuts) and acc >= total * cuts[nxt]:
                edges.append(i + 1); nxt += 1
        edges.append(n)
        edges = sorted(set(edges))
        self.e2e_chunks = []
        xo_end = self.x_off + [self.x_host.numel()]
        yo_end = self.off + [self.arena_len]
        for a, b in zip(edges[:-1], edges[1:]):
            sub = (b200.MulMatNode * (b - a))(*[self.nodes[i] for i in range(a, b)])
            self.e2e_chunks.append((sub, b - a, xo_end[a], xo_end[b], yo_end[a], yo_end[b]))
        self.s_h2d, self.s_d2h = torch.cuda.Stream(), torch.cuda.Stream()

    def step_e2e(self):
        """pinned host activations -> device, every MUL_MAT, outputs -> pinned host, through the C ABI.  Single GPU:
        pipelined over sub-lists (copies overlap the persistent launches).  Row split: the step is one fused launch; only
        rank 0 moves host data (every rank holds the gathered outputs; one copy of them leaves the box)."""
        torch = self.torch
        if self.world > 1:
            if self.rank == 0:
                self.x_dev.copy_(self.x_host, non_blocking=True)
            self.compute()
            if self.rank == 0:
                self.y_host[self.h0_len:].copy_(self.y_dev[self.h0_len:], non_blocking=True)
            return
        if not hasattr(self, "e2e_chunks"):
            self._e2e_setup(env_int("BENCH_E2E_CHUNKS", 4))
        cur = torch.cuda.current_stream()
        self.s_h2d.wait_stream(cur)                       # the previous step is done with x_dev
        self.s_d2h.wait_stream(cur)
        for sub, cnt, x0, x1, y0, y1 in self.e2e_chunks:
            with torch.cuda.stream(self.s_h2d):
                self.x_dev[x0:x1].copy_(self.x_host[x0:x1], non_blocking=True)
                ev_in = torch.cuda.Event(); ev_in.record(self.s_h2d)
            cur.wait_event(ev_in)
            self.kern.mul_mat_batch(sub, count=cnt, ws=self.ws)
            ev_out = torch.cuda.Event(); ev_out.record(cur)
            with torch.cuda.stream(self.s_d2h):
                self.s_d2h.wait_event(ev_out)
                self.y_host[y0:y1].copy_(self.y_dev[y0:y1], non_blocking=True)
        cur.wait_stream(self.s_d2h)                       # the step ends when its last output is on the host

    # ------------------------------------------------------------------ N-rank parity, carried by the bench line itself
    def selfcheck(self, sample_mats=4, sample_rows=256):
        """After a step every rank must hold every row.  Each rank regenerates the weights of ANOTHER rank's row block
        (same seeds), recomputes its first rows through the same kernel (a row's bits do not depend on the sub-matrix it
        is computed in) and compares with what the gather left in its own arena, bit for bit."""
        import b200
        torch, kern = self.torch, self.kern
        other = (self.rank + 1) % self.world
        checked, ok = 0, True
        step = max(1, len(self.mats) // sample_mats)
        for i in range(0, len(self.mats), step):
            mt = self.mats[i]
            lo, hi = self.bounds[i][other], self.bounds[i][other + 1]
            nr = min(sample_rows, hi - lo)
            w = random_blocks_device(torch, kern.device, mt.type, hi - lo, mt.k, self.weight_seed(i, other))
            rb = workloads.BLOCK_BYTES[mt.type] * (mt.k // workloads.BLOCK_ELEMS[mt.type])
            y = torch.zeros(nr, dtype=torch.float32, device=kern.device)
            one = (b200.MulMatNode * 1)(b200.mul_mat_node(mt.type, w.data_ptr(), nr, mt.k, self.x_dev.data_ptr() + 4 * self.x_off[i], 1, y.data_ptr()))
            kern.mul_mat_batch(one)
            got = self.y_dev[self.off[i] + lo: self.off[i] + lo + nr]
            ok = ok and bool(torch.equal(y.view(torch.int32), got.view(torch.int32))) and bool(torch.isfinite(y).all()) and bool(y.abs().max() > 0)
            checked += nr
            del w, rb
        return checked, ok


# ------------------------------------------------------------------------------------- timing
def timed_loop(torch, fn, steps, dist=None):
    """EXACTLY `steps` calls of fn between barrier + synchronize, CUDA events on the current stream;
    returns seconds (max over ranks)."""
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    start.record()
    for _ in range(steps):
        fn()
    stop.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    sec = start.elapsed_time(stop) * 1e-3
    if dist is not None:
        t = torch.tensor([sec], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item())
    return sec


def measure_per_node_launches(torch, kern, stack, reps):
    """Comparison leg: the same step issued as one launch per MUL_MAT (GGML_B200_NO_BATCH_KERNEL=1: the
    per-node fused decode kernel with programmatic dependent launch), replayed from a CUDA graph.
    Returns seconds per step."""
    prev = kern.set_option("no_batch_kernel", 1)
    try:
        stack.compute()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            stack.compute()
        g.replay()
        sec = timed_loop(torch, g.replay, reps) / reps
    finally:
        kern.set_option("no_batch_kernel", prev)
    return sec


def measure_prefill(torch, kern, n=512, reps=5):
    """Secondary figure (BASELINE configs[1]): batched MUL_MAT n=512 on the Llama-2-7B FFN shapes."""
    out = {}
    for mt in workloads.prefill_ffn_shapes():
        w = random_blocks_device(torch, kern.device, mt.type, mt.m, mt.k, 77)
        x = torch.rand((n, mt.k), device=kern.device) * 2 - 1
        y = torch.empty((n, mt.m), dtype=torch.float32, device=kern.device)
        fn = lambda: kern.mul_mat(mt.type, w, mt.m, mt.k, x, out=y)
        for _ in range(3):
            fn()
        sec = timed_loop(torch, fn, reps) / reps
        out[mt.name] = round(mt.flops(n) / sec / 1e12, 2)
        del w, x, y
    return out


def measure_support(torch, kern, peak):
    """Secondary figures: the support kernels of the path against the HBM peak;
    bytes = bytes read + bytes written.  tools/support_probe.py has the full format table."""
    import b200
    out = {}
    m, k = 32000, 4096
    w = random_blocks_device(torch, kern.device, "q4_K", m, k, 3)
    rows = torch.randint(0, m, (8192,), dtype=torch.int32, device=kern.device)
    x = torch.rand((4096, 8192), device=kern.device) * 2 - 1
    legs = {"dequantize_row_q4_K->f32": (lambda: kern.dequantize_row("q4_K", w, m * k, "f32"), w.numel() + m * k * 4),
            "get_rows_q4_K(8192 rows)": (lambda: kern.get_rows("q4_K", w, k, rows), 8192 * (b200.row_bytes("q4_K", k) + 4 * k)),
            "quantize_row_q8_1": (lambda: kern.quantize_row_q8_1(x), x.numel() * 4 + x.numel() // 32 * 36)}
    for name, (fn, byts) in legs.items():
        for _ in range(3):
            fn()
        sec = timed_loop(torch, fn, 10) / 10
        out[name] = {"GBps": round(byts / sec / 1e9, 1), "frac": round(byts / sec / 1e9 / peak, 3), "timed": "eager calls"}
        # a call is 36-100 us of GPU work behind a Python wrapper: replayed from a CUDA graph (3 calls per replay) the host gaps are gone
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                keep = [fn() for _ in range(3)]
            g.replay()
            sec_g = timed_loop(torch, g.replay, 10) / 30
            out[name] = {"GBps": round(byts / sec_g / 1e9, 1), "frac": round(byts / sec_g / 1e9 / peak, 3), "timed": "CUDA-graph replay",
                         "eager_GBps": out[name]["GBps"]}
            del g, keep
        except Exception:                        # noqa: BLE001 - the eager figure stands
            torch.cuda.synchronize()
    return out


def measure_gguf_load(torch, kern, gib=1.0):
    """Secondary figure: GGUF file (page cache, /tmp) -> HBM through b200_gguf_load (pinned staging ring + async H2D),
    wall clock around the call + stream sync; the device image is checked against the file afterwards."""
    import tempfile
    import time
    import numpy as np
    import b200
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import gguf_files as gf
    rb = b200.row_bytes("q4_K", 4096)
    n_t = max(2, int(gib * (1 << 30) / (11008 * rb)))
    rng = np.random.default_rng(1)
    raw = rng.integers(0, 256, size=11008 * rb, dtype=np.uint8).tobytes()
    tensors = [(f"blk.{i}.ffn_up.weight", [4096, 11008], 12, raw) for i in range(n_t)]
    with tempfile.TemporaryDirectory(dir="/tmp") as d:
        p = os.path.join(d, "bench.gguf")
        gf.write(p, [("general.architecture", gf.STR, "llama")], tensors)
        with b200.Gguf(p, kern.lib) as g:
            buf = torch.empty(g.device_bytes, dtype=torch.uint8, device=kern.device)
            best = None
            for _ in range(3):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                kern._check(kern.lib.b200_gguf_load(g.h, buf.data_ptr(), buf.numel(), kern._stream()), "gguf_load")
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
            last = g.tensors()[-1]
            ok = buf[last["device_offset"]:last["device_offset"] + last["nbytes"]].cpu().numpy().tobytes() == raw
            nbytes = g.data_bytes
    return {"GBps": round(nbytes / best / 1e9, 2), "file_GB": round(nbytes / 1e9, 3), "tensors": n_t, "image_matches_file": bool(ok),
            "path": "pread (page cache) -> 8 x 16 MiB pinned staging ring (one reader thread per stage) -> cudaMemcpyAsync"}


def measure_decode_batches(torch, kern, stack, peak):
    """Secondary figure: the same stack (same resident weights) with n = 2, 4, 8 activation columns per MUL_MAT
    (decode at batch n, BASELINE configs[3]): one b200_mul_mat_batch list per pass (the MMA list
    kernel of mmvq_cols.cuh), CUDA graph replay; GB/s on the weight bytes and tokens/s."""
    import b200
    out = {}
    mats = stack.mats
    wbytes = stack.local_weight_bytes
    for n in (2, 4, 8):
        xs = [torch.rand((n, mt.k), device=kern.device) * 2 - 1 for mt in mats]
        ys = [torch.empty((n, hi - lo), dtype=torch.float32, device=kern.device) for (lo, hi) in stack.rows]
        nodes = (b200.MulMatNode * len(mats))(*[
            b200.mul_mat_node(mt.type, w.data_ptr(), hi - lo, mt.k, x.data_ptr(), n, y.data_ptr(), const_weights=True)
            for mt, w, x, y, (lo, hi) in zip(mats, stack.weights, xs, ys, stack.rows)])
        ws = torch.empty(int(kern.lib.b200_mul_mat_batch_workspace_size(nodes, len(mats))) + 256, dtype=torch.uint8, device=kern.device)
        fn = lambda: kern.mul_mat_batch(nodes, ws=ws)    # noqa: E731
        fn()
        torch.cuda.synchronize()
        c0 = kern.launch_count()
        fn()
        launches = kern.launch_count() - c0
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        g.replay()
        sec = timed_loop(torch, g.replay, 10) / 10
        out[f"n={n}"] = {"ms_per_pass": round(sec * 1e3, 4), "launches": launches, "weight_GBps": round(wbytes / sec / 1e9, 1),
                         "frac": round(wbytes / sec / 1e9 / peak, 3), "tok_s": round(n / sec, 1)}
        del g, xs, ys, ws
    return out


def measure_decode_chain(torch, kern, stack, peak, reps=50, dist=None):
    """The same resident stack with the dependency structure a decoder pass really has (what a ggml decode graph gives
    the backend): per layer q/k/v read the layer input, wo reads the last of them (stand-in for the attention output),
    gate/up read wo, down reads up, the next layer reads down, the output matrix reads the last down -- 129 chain links
    for Llama-2-7B.  ONE list call: the dependent nodes ride the same persistent launch (completion counters +
    in-launch activation quantization; row split: the gathered vectors feed the next node inside the launch).
    Checked bit for bit against the same list cut at every link (option no_chain: one launch group per link, round 1's
    behaviour), which is also timed (single GPU)."""
    import b200
    stack.calibrate_chain()
    plan_rows = b200.batch_plan(kern.lib, stack.chain_nodes)
    out = {"nodes": len(stack.plan), "launch_groups": len({g for g, _, _ in plan_rows}),
           "in_launch_waits": sum(1 for _, dep, _ in plan_rows if dep >= 0),
           "structure": "per layer: first group reads the layer input, each later group reads the last output of the group before (q/k/v -> wo -> gate/up -> down -> next layer), output matrix last"}
    arena = stack.y_dev[stack.h0_len:]
    results = {}
    modes = ("chained", "cut") if stack.world == 1 else ("chained",)
    for mode in modes:
        prev = kern.set_option("no_chain", 1 if mode == "cut" else 0)
        try:
            arena.zero_()
            if dist is not None:
                torch.cuda.synchronize(); dist.barrier()
            stack.compute_chain()
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            results[mode] = arena.clone()
            c0 = kern.launch_count()
            stack.compute_chain()
            launches = kern.launch_count() - c0
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                stack.compute_chain()
            g.replay()
            sec = timed_loop(torch, g.replay, reps, dist) / reps
            alg = stack.local_alg_bytes * stack.world
            out[mode] = {"ms_per_pass": round(sec * 1e3, 4), "launches": launches, "GBps": round(alg / sec / 1e9, 1),
                         "frac_per_gpu": round(alg / stack.world / sec / 1e9 / peak, 3), "passes_per_s": round(1.0 / sec, 1)}
            torch.cuda.synchronize()
            del g
        finally:
            kern.set_option("no_chain", prev)
    if "cut" in results:
        out["bit_identical_to_cut_list"] = bool(torch.equal(results["chained"].view(torch.int32), results["cut"].view(torch.int32)))
    out["finite"] = bool(torch.isfinite(results["chained"]).all())
    return out


# ------------------------------------------------------------------------------------- end to end through the ggml plugin
class PluginStep:
    """The step as an unmodified ggml program runs it on our backend: libggml-b200.so loaded into the reference's own
    libggml (oracle/_ref/bin) with ggml_backend_load, a ggml graph of the workload's MUL_MAT nodes allocated on device
    B200_0, and per step
        ggml_backend_tensor_set_async(X  <- pinned host)   one tensor holding every activation vector (the nodes read views)
        ggml_backend_graph_compute_async                   the plugin hands the MUL_MAT run to b200_mul_mat_batch, CUDA-graph replay
        ggml_backend_tensor_get_async(Y  -> pinned host)   one tensor holding every output (placed with ggml_backend_tensor_alloc)
        ggml_backend_synchronize
    Pinned host memory comes from ggml_backend_b200_host_buffer_type.  chain=True builds the decoder-pass dependency
    structure instead (node outputs feed the next nodes directly; host traffic = first activation in, last output out)."""

    def __init__(self, stack, chain=False, chunks=1):
        import ctypes as C
        import numpy as np
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import ggml_host
        self.C, self.np = C, np
        ref_bin = os.path.join(ROOT, "oracle", "_ref", "bin")
        plugin = os.path.join(PKG, "libggml-b200.so")
        if not (os.path.exists(os.path.join(ref_bin, "libggml-base.so")) and os.path.exists(plugin)):
            raise RuntimeError("oracle/_ref or libggml-b200.so not built")
        host = self.host = ggml_host.GgmlHost(ref_bin)
        b = host.base
        vp, sz, i64 = C.c_void_p, C.c_size_t, C.c_int64
        b.ggml_view_2d.restype = vp;                   b.ggml_view_2d.argtypes = [vp, vp, i64, i64, sz, sz]
        b.ggml_get_data.restype = vp;                  b.ggml_get_data.argtypes = [vp]
        b.ggml_backend_tensor_alloc.restype = C.c_int; b.ggml_backend_tensor_alloc.argtypes = [vp, vp, vp]
        b.ggml_backend_tensor_set_async.restype = None; b.ggml_backend_tensor_set_async.argtypes = [vp, vp, vp, sz, sz]
        b.ggml_backend_tensor_get_async.restype = None; b.ggml_backend_tensor_get_async.argtypes = [vp, vp, vp, sz, sz]
        b.ggml_backend_buft_alloc_buffer.restype = vp; b.ggml_backend_buft_alloc_buffer.argtypes = [vp, sz]
        b.ggml_backend_buffer_get_base.restype = vp;   b.ggml_backend_buffer_get_base.argtypes = [vp]
        b.ggml_backend_buffer_set_usage.restype = None; b.ggml_backend_buffer_set_usage.argtypes = [vp, C.c_int]
        b.ggml_backend_event_new.restype = vp;         b.ggml_backend_event_new.argtypes = [vp]
        b.ggml_backend_event_free.restype = None;      b.ggml_backend_event_free.argtypes = [vp]
        b.ggml_backend_event_record.restype = None;    b.ggml_backend_event_record.argtypes = [vp, vp]
        b.ggml_backend_event_wait.restype = None;      b.ggml_backend_event_wait.argtypes = [vp, vp]
        host.load_plugin(plugin)
        dev = host.device_by_name("B200_0")
        if not dev:
            raise RuntimeError("device B200_0 not registered")
        self.backend = host.init_device(dev)
        plug = C.CDLL(plugin)
        plug.ggml_backend_b200_host_buffer_type.restype = vp
        mats = stack.mats
        n = len(mats)
        T = ggml_host.GGML_TYPE
        # weights: their own context and buffer, marked as weights (what llama.cpp does), uploaded once
        self.ctx_w = host.new_context(n_tensors=n + 8)
        self.w = [b.ggml_new_tensor_2d(self.ctx_w, T[mt.type], mt.k, hi - lo) for mt, (lo, hi) in zip(mats, stack.rows)]
        self.buf_w = b.ggml_backend_alloc_ctx_tensors(self.ctx_w, self.backend)
        if not self.buf_w:
            raise MemoryError("weights")
        b.ggml_backend_buffer_set_usage(self.buf_w, 1)           # GGML_BACKEND_BUFFER_USAGE_WEIGHTS
        for t, wt in zip(self.w, stack.weights):
            host.tensor_set(t, wt.cpu().numpy())
        # activations X (every vector, the nodes read views) and outputs Y (every result, placed by hand) in one buffer
        plan = stack.plan if chain else [(i, None, None) for i in range(n)]
        xlen = stack.h0_len if chain else stack.x_host.numel()
        ylen = sum((mt.m + 31) // 32 * 32 for mt in mats)
        self.ctx_io = host.new_context(n_tensors=2 * n + 16, graph_nodes=2 * n + 64)
        self.X = b.ggml_new_tensor_1d(self.ctx_io, T["f32"], xlen)
        self.Y = b.ggml_new_tensor_1d(self.ctx_io, T["f32"], ylen)
        xv = []
        if not chain:
            for i, mt in enumerate(mats):
                xv.append(b.ggml_view_2d(self.ctx_io, self.X, mt.k, 1, 4 * mt.k, 4 * stack.x_off[i]))
        else:
            xv.append(b.ggml_view_2d(self.ctx_io, self.X, stack.h0_len, 1, 4 * stack.h0_len, 0))
        self.buf_io = b.ggml_backend_alloc_ctx_tensors(self.ctx_io, self.backend)
        if not self.buf_io:
            raise MemoryError("activations")
        self.ctx_g = host.new_context(n_tensors=n + 16, graph_nodes=2 * n + 64)
        ybase = b.ggml_get_data(self.Y)
        ys, yoff, pos = {}, [], 0
        for idx, (i, src, _) in enumerate(plan):
            mt = mats[i]
            x = xv[i] if not chain else (xv[0] if src == "h0" else ys[src])
            y = b.ggml_mul_mat(self.ctx_g, self.w[i], x)
            if b.ggml_backend_tensor_alloc(self.buf_io, y, ybase + 4 * pos) != 0:
                raise RuntimeError("ggml_backend_tensor_alloc")
            ys[f"y{i}"] = y
            yoff.append(pos)
            pos += (mt.m + 31) // 32 * 32
        # `chunks` > 1 (independent list only): the step as `chunks` sub-graphs over consecutive node ranges, so that a program
        # with three backend instances on the device (= three streams) can overlap the host copies of one range with the
        # compute of another, ordered by ggml_backend_event_record / _wait -- what ggml_backend_sched does between splits
        if chain:
            chunks = 1
        edges = [0, n]
        if chunks > 1:
            total = sum(int(wt.numel()) for wt in stack.weights)
            edge = 0.1
            cuts = [edge + (1.0 - 2.0 * edge) * j / (chunks - 2) for j in range(chunks - 1)] if chunks >= 4 else [j / chunks for j in range(1, chunks)]
            edges, acc, nxt = [0], 0, 0
            for i, wt in enumerate(stack.weights):
                acc += int(wt.numel())
                while nxt < len(cuts) and acc >= total * cuts[nxt]:
                    edges.append(i + 1); nxt += 1
            edges = sorted(set(edges + [n]))
        self.ranges, self.graphs = [], []
        last = None
        for a_, b_ in zip(edges[:-1], edges[1:]):
            ctx_c = host.new_context(n_tensors=8, graph_nodes=2 * (b_ - a_) + 64)
            g = b.ggml_new_graph_custom(ctx_c, max(256, 2 * (b_ - a_) + 64), False)
            for idx in range(a_, b_):
                b.ggml_build_forward_expand(g, ys[f"y{plan[idx][0]}"])
                last = plan[idx][0]
            x0 = 0 if chain else 4 * stack.x_off[a_]
            x1 = 4 * xlen if (chain or b_ == n) else 4 * stack.x_off[b_]
            y0 = 4 * yoff[a_]
            y1 = 4 * ylen if b_ == n else 4 * yoff[b_]
            self.graphs.append((ctx_c, g))
            self.ranges.append((x0, x1, y0, y1))
        self.graph = self.graphs[0][1]
        self.chain, self.yoff, self.ys, self.last, self.chunks = chain, yoff, ys, last, len(self.graphs)
        self.be_in = self.be_out = None
        if self.chunks > 1:
            self.be_in, self.be_out = host.init_device(dev), host.init_device(dev)
            self.ev_in = [b.ggml_backend_event_new(dev) for _ in self.graphs]
            self.ev_out = [b.ggml_backend_event_new(dev) for _ in self.graphs]
        # pinned host staging from the plugin's host buffer type
        buft = plug.ggml_backend_b200_host_buffer_type()
        self.h_bytes_in = 4 * xlen
        self.h_bytes_out = 4 * ylen if not chain else 4 * mats[last].m
        self.hbuf = b.ggml_backend_buft_alloc_buffer(buft, self.h_bytes_in + 4 * ylen + 256)
        if not self.hbuf:
            raise MemoryError("pinned host buffer")
        hbase = b.ggml_backend_buffer_get_base(self.hbuf)
        self.h_in, self.h_out = hbase, hbase + (self.h_bytes_in + 127) // 128 * 128
        src_host = stack.y_dev[:stack.h0_len].cpu().numpy() if chain else stack.x_host.numpy()
        C.memmove(self.h_in, src_host.ctypes.data, self.h_bytes_in)

    def step(self):
        b = self.host.base
        if self.chunks > 1:
            for c, ((x0, x1, y0, y1), (_, g)) in enumerate(zip(self.ranges, self.graphs)):
                b.ggml_backend_tensor_set_async(self.be_in, self.X, self.h_in + x0, x0, x1 - x0)
                b.ggml_backend_event_record(self.ev_in[c], self.be_in)
                b.ggml_backend_event_wait(self.backend, self.ev_in[c])
                if b.ggml_backend_graph_compute_async(self.backend, g) != 0:
                    raise RuntimeError("graph_compute")
                b.ggml_backend_event_record(self.ev_out[c], self.backend)
                b.ggml_backend_event_wait(self.be_out, self.ev_out[c])
                b.ggml_backend_tensor_get_async(self.be_out, self.Y, self.h_out + y0, y0, y1 - y0)
            b.ggml_backend_synchronize(self.be_out)
            b.ggml_backend_synchronize(self.backend)
            b.ggml_backend_synchronize(self.be_in)
            return
        b.ggml_backend_tensor_set_async(self.backend, self.X, self.h_in, 0, self.h_bytes_in)
        st = b.ggml_backend_graph_compute_async(self.backend, self.graph)
        if st != 0:
            raise RuntimeError(f"graph_compute status {st}")
        if self.chain:
            b.ggml_backend_tensor_get_async(self.backend, self.ys[f"y{self.last}"], self.h_out, 0, self.h_bytes_out)
        else:
            b.ggml_backend_tensor_get_async(self.backend, self.Y, self.h_out, 0, self.h_bytes_out)
        b.ggml_backend_synchronize(self.backend)

    def outputs(self):
        """host copy of the last step's outputs as a float32 array (all outputs, or the chain's last vector)"""
        n = self.h_bytes_out // 4
        return self.np.ctypeslib.as_array((self.C.c_float * n).from_address(self.h_out)).copy()

    def timed(self, steps, warmup=3):
        for _ in range(max(3, warmup)):                  # the plugin captures a repeated graph on its 2nd sight, replays from the 3rd
            self.step()
        t0 = time.perf_counter()
        for _ in range(steps):
            self.step()
        return time.perf_counter() - t0

    def close(self):
        b = self.host.base
        for buf in (self.buf_io, self.buf_w, self.hbuf):
            if buf:
                b.ggml_backend_buffer_free(buf)
        for ctx_c, _ in self.graphs:
            b.ggml_free(ctx_c)
        b.ggml_free(self.ctx_g); b.ggml_free(self.ctx_io); b.ggml_free(self.ctx_w)
        if self.chunks > 1:
            for ev in self.ev_in + self.ev_out:
                b.ggml_backend_event_free(ev)
            b.ggml_backend_free(self.be_in); b.ggml_backend_free(self.be_out)
        b.ggml_backend_free(self.backend)


# ------------------------------------------------------------------------------------- the CPU reference
def cpu_reference(budget_s=12.0, sample_layers=4, steps=None, warmup=1, stack_name="llama2-7b"):
    """The reference's own CPU implementation of the path (ggml-cpu, unmodified, compiled from
    /root/reference into oracle/_ref by oracle/Makefile) on a bounded sample of the workload:
    `sample_layers` Llama-2-7B layers (7 matrices each, Q4_K, n = 1) as one ggml graph on the CPU
    backend with every host thread.  Returns (GB/s on algorithmic bytes, dict)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ggml_host
    import oracle_lib
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "bin")
    cores = os.cpu_count() or 1
    if stack_name == "llama3-70b-ffn":
        mats = workloads.llama3_70b_ffn(n_layers=1)                 # 3 matrices, 397 MB >> host L3
    else:
        mats = workloads.llama2_7b_stack("q4_K", n_layers=sample_layers)[:-1]
    alg = sum(mt.algorithmic_bytes(1) for mt in mats)
    if os.path.exists(os.path.join(ref_bin, "libggml-base.so")):
        host = ggml_host.GgmlHost(ref_bin)
        cpu = host.cpu_backend(n_threads=cores)
        g = ggml_host.MulMatGraph(host, cpu, mats, 1)
        rng = np.random.default_rng(5)
        protos = {}
        for i, mt in enumerate(mats):
            if (mt.m, mt.k) not in protos:
                protos[(mt.m, mt.k)] = oracle_lib.random_blocks(mt.type, mt.m, mt.k, seed=i)
            g.set_weight(i, protos[(mt.m, mt.k)])
            g.set_input(i, rng.uniform(-1, 1, size=(1, mt.k)).astype(np.float32))
        for _ in range(max(1, warmup)):               # warm-up (thread pool start, page faults)
            g.compute()
        t0 = time.perf_counter(); reps = 0
        while True:
            g.compute(); reps += 1
            el = time.perf_counter() - t0
            if (steps is not None and reps >= steps) or (steps is None and (el >= budget_s or reps >= 2000)):
                break
        g.close()
        kind = "reference"
    else:
        # the C restatement (scalar, 1 thread) when the reference build products are absent
        orc = oracle_lib.Oracle()
        cores = 1
        mats = mats[:7]
        alg = sum(mt.algorithmic_bytes(1) for mt in mats)
        ws = [oracle_lib.random_blocks(mt.type, mt.m, mt.k, seed=i) for i, mt in enumerate(mats)]
        xs = [np.random.default_rng(i).uniform(-1, 1, size=(1, mt.k)).astype(np.float32) for i, mt in enumerate(mats)]
        t0 = time.perf_counter(); reps = 0
        while True:
            for mt, w, x in zip(mats, ws, xs):
                orc.mul_mat(mt.type, w, mt.m, mt.k, x, mode=1)
            reps += 1
            el = time.perf_counter() - t0
            if el >= budget_s:
                break
        kind = "port"
    gbps = alg * reps / el / 1e9
    return gbps, {"value": round(gbps, 2), "unit": UNIT, "cores": cores, "kind": kind,
                  "sample": (f"{len(mats)} matrices (1 Llama-3-70B FFN layer, Q4_K, n=1, " if stack_name == "llama3-70b-ffn" else
                             f"{len(mats)} matrices ({len(mats) // 7} Llama-2-7B layers, Q4_K, n=1, ")
                            + f"{alg / 1e6:.0f} MB) x {reps} passes in {el:.1f} s",
                  "ms_per_pass": round(el / reps * 1e3, 3)}


# ------------------------------------------------------------------------------------- main
def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """dram bytes (read + write) per launch of the dominant kernel from the committed ncu --set full
    capture (profiles/traffic.json; the capture is of this same workload)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return int(json.load(f)["dram_bytes_per_launch"])
    except (OSError, ValueError, KeyError):
        return None


def run_reference(args, rank):
    if rank != 0:
        return
    stack_name = args.stack or ("llama2-7b" if args.gpus <= 1 else "llama3-70b-ffn")
    gbps, info = cpu_reference(steps=args.steps, warmup=args.warmup, stack_name=stack_name)
    wl = WORKLOAD if stack_name == "llama2-7b" else "llama3-70b FFN (gate, up 28672x8192; down 8192x28672), Q4_K, decode mat-vec n=1 (BASELINE configs[4])"
    line = {"impl": "reference", "metric": METRIC, "value": info["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": info["ms_per_pass"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "int8 dot (q8_K activations), f32 accumulate",
            "data": "synthetic", "config": {"workload": wl, "n": 1, "sample": info["sample"]},
            "cpu_baseline": {k: info[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": info["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def tensor_peaks(kern):
    """MEASURED_PEAKS.json (driver: cuBLAS bf16 8192^3, burst and sustained) + this run's tcgen05 micro-kernel
    (b200_tensor_peak: operands resident in shared memory, nothing but the MMA issue loop) for bf16 and int8."""
    out = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mp = json.load(f)
        out["bf16_cublas_burst_TFLOPs"] = mp.get("bf16_tflops"); out["bf16_cublas_sustained_TFLOPs"] = mp.get("bf16_tflops_sustained")
    except (OSError, ValueError):
        out["bf16_cublas_burst_TFLOPs"], out["bf16_cublas_sustained_TFLOPs"] = 1590.0, 1400.0
        out["note"] = "MEASURED_PEAKS.json absent: fallback figures of B200_PROFILING.md"
    try:
        out["tcgen05_bf16_TFLOPs"] = round(kern.tensor_peak("bf16", iters=8192, reps=5) / 1e12, 1)
        out["tcgen05_int8_TOPs"] = round(kern.tensor_peak("i8", iters=8192, reps=5) / 1e12, 1)
    except Exception as e:                       # noqa: BLE001
        out["error"] = str(e)[:200]
    return out


def measure_prefill_all(torch, kern, peaks):
    """The batched (tensor-core) term of the metric: BASELINE configs[1] (Q8_0 / Q4_0, n = 512, Llama-2-7B FFN shapes) and
    configs[3] prefill shapes at n = 2048, TFLOP/s = 2 m n k / time, with the roofline of the best figure."""
    out = {"n512": measure_prefill(torch, kern, n=512)}
    n2048 = {}
    for mt in (workloads.Mat("q4_K.14336x4096", "q4_K", 14336, 4096), workloads.Mat("q6_K.4096x14336", "q6_K", 4096, 14336),
               workloads.Mat("q8_0.11008x4096", "q8_0", 11008, 4096)):
        w = random_blocks_device(torch, kern.device, mt.type, mt.m, mt.k, 78)
        x = torch.rand((2048, mt.k), device=kern.device) * 2 - 1
        y = torch.empty((2048, mt.m), dtype=torch.float32, device=kern.device)
        fn = lambda: kern.mul_mat(mt.type, w, mt.m, mt.k, x, out=y)      # noqa: E731
        for _ in range(3):
            fn()
        sec = timed_loop(torch, fn, 5) / 5
        n2048[mt.name] = round(mt.flops(2048) / sec / 1e12, 2)
        del w, x, y
    out["n2048"] = n2048
    best = max(list(out["n512"].values()) + list(n2048.values()))
    sustained = peaks.get("bf16_cublas_sustained_TFLOPs") or 1400.0
    out["roofline"] = {"bound": "tensor", "kernel": "k_mmq_tc (tcgen05 kind::f16, bf16 operands, f32 TMEM accumulator)", "achieved": best,
                       "peak": sustained, "unit": "TFLOP/s", "frac": round(best / sustained, 4),
                       "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (cuBLAS, driver-measured)",
                       "frac_of_tcgen05_bf16_micro_kernel": round(best / peaks["tcgen05_bf16_TFLOPs"], 4) if peaks.get("tcgen05_bf16_TFLOPs") else None,
                       "frac_of_tcgen05_int8_micro_kernel": round(best / peaks["tcgen05_int8_TOPs"], 4) if peaks.get("tcgen05_int8_TOPs") else None,
                       "traffic": None}
    return out


def capture(torch, fn):
    """fn captured into a CUDA graph (None when capture fails: the step then runs eagerly)"""
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        return g
    except Exception as e:                       # noqa: BLE001 - capture is an optimisation, not a requirement
        sys.stderr.write(f"[bench] CUDA graph capture failed ({e}); timing eager launches\n")
        torch.cuda.synchronize()
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--wtype", default="q4_K", help="weight format of the llama2-7b stack (default: the headline Q4_K)")
    ap.add_argument("--stack", default=None, choices=["llama2-7b", "llama3-70b-ffn"],
                    help="default: llama2-7b at 1 GPU (BASELINE configs[2]); llama3-70b-ffn at N > 1 (configs[4]: 16 layers of the "
                         "8192x28672 Q4_K FFN, rows split over the N GPUs, strong scaling).  llama2-7b at N > 1: rows x N (weak scaling)")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary figures and the CPU baseline")
    args = ap.parse_args()
    rank = env_int("RANK", 0)
    world = env_int("WORLD_SIZE", 1)
    local_rank = env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)
    if args.stack is None:
        args.stack = "llama2-7b" if world == 1 else "llama3-70b-ffn"

    import torch
    import b200
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dbg("process group up")
    kern = b200.Kernels(device=local_rank)
    strong = args.stack == "llama3-70b-ffn"
    mats = workloads.llama3_70b_ffn(n_layers=16) if strong else workloads.llama2_7b_stack(args.wtype, row_mult=world)
    stack = DecodeStack(kern, mats, rank, world, dist)
    lib = kern.lib
    fused = world > 1 and stack.region is not None
    dbg("stack built")

    # ---- warm-up (eager), then capture the step into a CUDA graph
    for _ in range(args.warmup):
        stack.compute()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    dbg("eager warm-up done")
    c0 = int(lib.b200_launch_count())
    stack.compute()
    launches_per_step = int(lib.b200_launch_count()) - c0
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    graph = capture(torch, stack.compute)
    step = graph.replay if graph is not None else stack.compute
    dbg(f"capture done (graph={graph is not None})")
    for _ in range(3):
        step()
    torch.cuda.synchronize()

    # ---- device-resident timed region
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    sec = timed_loop(torch, step, args.steps, dist)
    dbg("device-resident region done")

    # ---- N-rank parity carried by the line itself, and the NCCL baseline leg
    multi = None
    if world > 1:
        torch.cuda.synchronize(); dist.barrier()
        checked, ok = stack.selfcheck()
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        multi = {"gather": "fused into the persistent launch (peer stores over NVLink + system-scope completion flags)" if fused
                           else "NCCL all_gather_into_tensor per MUL_MAT group (symmetric memory unavailable)",
                 "selfcheck": {"rows_recomputed_per_rank": checked, "of_rank": "(rank + 1) % N", "bit_identical_on_every_rank": bool(flag.item())}}
        if fused and not args.no_extras:
            try:
                for _ in range(3):
                    stack.compute_nccl()
                torch.cuda.synchronize(); dist.barrier()
                c0 = int(lib.b200_launch_count())
                stack.compute_nccl()
                nl = int(lib.b200_launch_count()) - c0
                gn = capture(torch, stack.compute_nccl)
                fn = gn.replay if gn is not None else stack.compute_nccl
                for _ in range(3):
                    fn()
                reps = max(10, min(args.steps, 50))
                sec_n = timed_loop(torch, fn, reps, dist) / reps
                multi["nccl_per_group_baseline"] = {"ms_per_step": round(sec_n * 1e3, 4), "groups": len(stack.groups), "launches": nl,
                                                    "GBps": round(stack.local_alg_bytes * world / sec_n / 1e9, 1)}
                del gn
            except Exception as e:               # noqa: BLE001
                multi["nccl_per_group_baseline"] = {"error": str(e)[:200]}
        dbg("multi-rank legs done")

    # ---- end to end, host buffers, every step
    e2e_path, plug = None, None
    if world == 1:
        try:
            plug = PluginStep(stack, chunks=env_int("BENCH_E2E_CHUNKS", 4))
            sec_e2e = plug.timed(args.steps, args.warmup)
            out = plug.outputs()
            ref_out = stack.y_dev[stack.h0_len:].cpu().numpy()
            if out.shape != ref_out.shape or not (out.view("uint32") == ref_out.view("uint32")).all():
                raise RuntimeError("plugin outputs differ from the C-ABI outputs")
            e2e_path = ("pinned host (ggml_backend_b200_host_buffer_type) -> ggml_backend_tensor_set_async -> ggml_backend_graph_compute_async "
                        "on B200_0 (plugin: b200_mul_mat_batch, CUDA-graph replay) -> ggml_backend_tensor_get_async -> ggml_backend_synchronize; "
                        + (f"the step as {plug.chunks} sub-graphs on three backend instances of the device (copy-in / compute / copy-out streams "
                           "ordered by ggml_backend_event_record / _wait), so host copies overlap compute; " if plug.chunks > 1 else "")
                        + "outputs bit-identical to the C-ABI step; host wall clock")
            h2d, d2h = plug.h_bytes_in, plug.h_bytes_out
        except Exception as e:                   # noqa: BLE001
            sys.stderr.write(f"[bench] plugin e2e leg unavailable ({str(e)[:300]}); timing the C-ABI e2e\n")
            plug = None
    if plug is None:
        def e2e_step():
            stack.step_e2e()
            torch.cuda.current_stream().synchronize()
        for _ in range(3):
            e2e_step()
        sec_e2e = timed_loop(torch, e2e_step, args.steps, dist)
        e2e_path = ("pinned host activations -> b200_mul_mat_batch" + ("_peers" if fused else "") + " (C ABI) -> pinned host outputs, sync per step"
                    + ("; rank 0 moves the host data (every rank holds the gathered outputs)" if world > 1 else "; pipelined over 4 sub-lists"))
        h2d, d2h = stack.x_bytes, stack.y_bytes
    dbg("e2e region done")
    clk = clocks.stop() if rank == 0 else None
    if not bool(torch.isfinite(stack.y_dev).all()):
        raise RuntimeError("non-finite outputs")

    alg_total = stack.local_alg_bytes * world                       # every rank does the same amount
    value = alg_total * args.steps / sec / 1e9
    e2e = alg_total * args.steps / sec_e2e / 1e9
    weight_total = stack.local_weight_bytes * world

    # ---- roofline of the dominant kernel, measured live: a step is one k_quantize_batch launch (a few us)
    # + one k_mmvq_batch launch, so the persistent kernel's launch duration is the step's device time (the
    # CUDA events of the timed region above, on the launching stream); its algorithmic bytes are the step's.
    k_sec, k_bytes = sec / args.steps, stack.local_alg_bytes
    peak, peak_src = hbm_peak()
    achieved = k_bytes / k_sec / 1e9
    traffic = ncu_traffic() if (world == 1 and args.stack == "llama2-7b" and args.wtype == "q4_K") else None
    wl = ("llama3-70b FFN (gate, up 28672x8192; down 8192x28672) x 16 layers, Q4_K, decode mat-vec n=1, rows split over the GPUs (BASELINE configs[4])"
          if strong else (WORKLOAD if args.wtype == "q4_K" else WORKLOAD.replace("Q4_K", args.wtype)))
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(sec / args.steps * 1e3, 4), "higher_is_better": True,
        "scaling": "strong" if (strong or world == 1) else "weak", "vs_baseline": None, "dtype": "int8 dot (q8_1 activations), f32 accumulate",
        "data": "synthetic",
        "config": {"workload": wl, "n": 1, "matrices": len(mats), "row_mult": 1 if strong else world,
                   "weight_bytes_per_gpu": stack.local_weight_bytes, "algorithmic_bytes_per_step": alg_total,
                   "l2": f"inputs larger than L2 ({stack.local_weight_bytes / 1e9:.2f} GB of weights per GPU streamed once per step; no flush)",
                   "launch": "CUDA graph replay" if graph is not None else "eager",
                   "parallelism": "single GPU" if world == 1 else
                                  (f"row-split x{world}: one persistent launch per rank; finished rows stored into every rank's arena from inside the launch" if fused
                                   else f"row-split x{world}: one persistent launch + one NCCL all-gather per MUL_MAT group")},
        "e2e": {"value": round(e2e, 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": round(sec_e2e / args.steps * 1e3, 4), "path": e2e_path},
        "gpu_launches": launches_per_step * args.steps,
        "weight_GBps": round(weight_total * args.steps / sec / 1e9, 1),
        "roofline": {"bound": "hbm", "kernel": "k_mmvq_batch (persistent list mat-vec; + k_quantize_batch pre-pass in the same timed launch pair)",
                     "achieved": round(achieved, 1),
                     "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4), "peak_source": peak_src,
                     "frac_of_8TBps_spec": round(achieved / 8000.0, 4),
                     "us_per_launch": round(k_sec * 1e6, 3), "bytes_per_launch": int(k_bytes),
                     "traffic": traffic},
        "clocks": clk,
    }
    if multi is not None:
        line["multi_gpu"] = multi
    if not args.no_extras:
        try:
            line["decode_chain"] = measure_decode_chain(torch, kern, stack, peak, reps=50, dist=dist)
            ch = line["decode_chain"].get("chained")
            if ch and world == 1:
                line["roofline_decode_tok_s"] = {"bound": "hbm", "what": "GEMM-only decode passes/s of the dependent chain (the tok/s term of the metric; no attention / norms)",
                                                 "achieved": ch["passes_per_s"], "peak": round(peak * 1e9 / stack.local_alg_bytes, 1), "unit": "passes/s",
                                                 "frac": round(ch["passes_per_s"] / (peak * 1e9 / stack.local_alg_bytes), 4), "traffic": None}
        except Exception as e:                   # noqa: BLE001
            line["decode_chain"] = {"error": str(e)[:300]}
    if plug is not None and not args.no_extras:
        try:
            pc = PluginStep(stack, chain=True)
            sec_c = pc.timed(min(args.steps, 100), 3)
            n_c = min(args.steps, 100)
            got = pc.outputs()
            li = stack.plan[-1][0]
            want = stack.y_dev[stack.off[li]: stack.off[li] + mats[li].m].cpu().numpy()
            line["decode_chain"]["e2e_through_plugin"] = {
                "ms_per_pass": round(sec_c / n_c * 1e3, 4), "GBps": round(stack.local_alg_bytes * n_c / sec_c / 1e9, 1),
                "h2d_bytes_per_step": pc.h_bytes_in, "d2h_bytes_per_step": pc.h_bytes_out,
                "logits_bit_identical_to_c_abi_chain": bool((got.view("uint32") == want.view("uint32")).all()),
                "path": "ggml graph with the chain's dependencies on B200_0: first activation in, last output out"}
            pc.close()
        except Exception as e:                   # noqa: BLE001
            line.setdefault("decode_chain", {})["e2e_through_plugin"] = {"error": str(e)[:300]}
    if plug is not None:
        plug.close()
    if rank == 0 and world == 1 and not args.no_extras:
        def leg(name, fn):
            try:
                line[name] = fn()
            except Exception as e:               # noqa: BLE001
                line[name] = {"error": str(e)[:200]}
        leg("one_launch_per_mul_mat", lambda: (lambda t: {"ms_per_step": round(t * 1e3, 4), "GBps": round(stack.local_alg_bytes / t / 1e9, 1),
                                                          "note": "same step as 225 per-node fused decode launches (CUDA graph + PDL), for comparison"})(
            measure_per_node_launches(torch, kern, stack, reps=max(5, min(args.steps, 20)))))
        peaks = tensor_peaks(kern)
        line["tensor_peaks"] = peaks
        leg("prefill_TFLOPs", lambda: measure_prefill_all(torch, kern, peaks))
        leg("support_kernels", lambda: measure_support(torch, kern, peak))
        leg("decode_batches", lambda: measure_decode_batches(torch, kern, stack, peak))
        leg("gguf_load", lambda: measure_gguf_load(torch, kern))
        def c_abi_e2e():
            def e2e_step():
                stack.step_e2e()
                torch.cuda.current_stream().synchronize()
            for _ in range(3):
                e2e_step()
            n_ = min(args.steps, 100)
            t = timed_loop(torch, e2e_step, n_) / n_
            return {"ms_per_step": round(t * 1e3, 4), "GBps": round(stack.local_alg_bytes / t / 1e9, 1),
                    "path": "b200_mul_mat_batch (C ABI) with torch pinned tensors, copies pipelined over 4 sub-lists on 3 streams"}
        leg("e2e_c_abi_pipelined", c_abi_e2e)
        _, info = cpu_reference()
        line["cpu_baseline"] = {k: info[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if rank == 0:
        print(json.dumps(line), flush=True)
    dbg("line printed")
    if dist is not None:
        # Tearing the process group down while a captured graph still holds NCCL work hung both ranks
        # (measured, NCCL 2.28.9): drop the graph, drain, meet once more and leave without the teardown.
        graph = None
        step = None
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        dbg("exit")
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
