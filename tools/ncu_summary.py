"""Summarise an ncu report (read here, no GPU): headline raw metrics + stall reasons + hottest SASS lines.
usage: ncu_summary.py REPORT.ncu-rep [top] [result index in the report]"""
import csv, subprocess, sys, io
from collections import defaultdict
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
sel = ["--launch-skip", sys.argv[3], "--launch-count", "1"] if len(sys.argv) > 3 else []
raw = subprocess.run(["ncu", "-i", rep] + sel + ["--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u, v = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max", "launch__shared_mem_per_block_dynamic",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "smsp__inst_executed_pipe_tensor.sum", "sm__inst_executed_pipe_uniform.sum",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "nvltx__bytes.sum", "nvlrx__bytes.sum"]
for i, k in enumerate(h):
    if k in want or "pipe_tensor" in k:
        print(f"{k:70s} {v[i]} {u[i]}")
src = subprocess.run(["ncu", "-i", rep] + sel + ["--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h, data = rows[1], rows[2:]
ix = {k: i for i, k in enumerate(h)}
def f(r, k):
    try: return float(r[ix[k]])
    except (ValueError, KeyError, IndexError): return 0.0
tot = sum(f(r, "# Samples") for r in data)
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
print(f"samples {tot:.0f}, instructions {sum(f(r, 'Instructions Executed') for r in data):.0f}")
for k, val in sorted(((k, sum(f(r, k) for r in data)) for k in stalls), key=lambda x: -x[1])[:8]:
    print(f"  {k:28s} {val / tot * 100:5.1f}%")
g = defaultdict(lambda: [0, 0.0])
for r in data:
    e = int(f(r, "Instructions Executed")); g[e][0] += 1; g[e][1] += f(r, "# Samples")
print("by execution count (groups of SASS lines executed equally often):")
for e, (n, sm) in sorted(g.items(), key=lambda x: -x[1][1])[:8]:
    print(f"  executed {e:9d} x {n:4d} lines: {sm / tot * 100:5.1f}% of samples")
print("hottest lines:")
for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:top]:
    t2 = sorted(stalls, key=lambda k: -f(r, k))[:2]
    print(f"  {f(r, '# Samples') / tot * 100:5.2f}%  x{int(f(r, 'Instructions Executed')):9d}  {r[ix['Source']][:64]:64s} {[(k[6:], int(f(r, k))) for k in t2]}")
