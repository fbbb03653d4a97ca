#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "dequant or get_rows or cpy or quantize" 2>&1 | tail -4 )
timeout 400 python tools/support_probe.py 2>&1 | tail -18 | tee gpurun_out/support_probe.log
