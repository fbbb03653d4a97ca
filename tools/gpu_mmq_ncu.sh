#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 120 python tools/mmq_once.py q4_0 11008 4096 2048 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmq_tc -s 1 -c 1 -o gpurun_out/r01l_mmq python tools/mmq_once.py q4_0 11008 4096 2048 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
