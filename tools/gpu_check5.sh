#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "persistent or batch or fused" 2>&1 | tail -3 )
timeout 300 python tools/shape_probe.py q4_K 2>&1 | tail -6 | tee gpurun_out/shape_probe.txt
for t in q4_K l3mix; do timeout 120 python tools/batch_probe.py 20 $t 2>&1 | tail -1; done | tee gpurun_out/batch_probe_types.log
