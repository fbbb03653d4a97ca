#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "persistent or batch or fused" 2>&1 | tail -15 ) > gpurun_out/pytest_batch.log 2>&1
tail -3 gpurun_out/pytest_batch.log
for t in q4_0 q8_0; do timeout 120 python tools/batch_probe.py 20 $t 2>&1 | tail -1; done | tee gpurun_out/batch_probe_types.log
