#!/bin/bash
# ncu --set full capture of the persistent kernel on the bench workload: gpu_ncu.sh NAME
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 120 python tools/batch_probe.py 0 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_batch -s 1 -c 1 -o gpurun_out/$1 python tools/batch_probe.py 0 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
