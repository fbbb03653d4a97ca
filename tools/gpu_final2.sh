#!/bin/bash
# 2-GPU sanity at round end: the split-buffer plugin tests (skipped on one GPU), smoke, and the 2-rank bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_parity.py -x -q -m gpu -k "row_split or few_columns or decode_graph" 2>&1 | tail -3
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
bash tools/gpu_multi.sh 2 100 2>&1 | tail -3 | cut -c1-400
