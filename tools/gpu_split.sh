#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_plugin.py -m gpu -x -q -k "split or decode_graph or graph_mul_mat" 2>&1 | tail -15 ) | tee gpurun_out/pytest_split.log
