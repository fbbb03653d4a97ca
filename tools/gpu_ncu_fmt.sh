#!/bin/bash
# full ncu capture of the persistent kernel on a one-format stack.  usage: gpu_ncu_fmt.sh TAG WTYPE
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
TAG=${1:-r01}; WT=${2:-q4_0}
mkdir -p gpurun_out
timeout 120 python tools/batch_probe.py 5 $WT 2>&1 | tail -1
timeout 120 python tools/batch_probe.py 0 $WT > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_batch -s 1 -c 1 -o gpurun_out/${TAG}_batch_$WT python tools/batch_probe.py 0 $WT > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
