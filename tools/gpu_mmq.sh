#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mul_mat_q" 2>&1 | tail -4 )
timeout 200 python tools/prefill_probe.py 2>&1 | tail -3 | tee gpurun_out/prefill_probe.log
GGML_B200_MMQ_BN=256 timeout 200 python tools/prefill_probe.py 2>&1 | tail -3 | tee -a gpurun_out/prefill_probe.log
