#!/bin/bash
# full ncu capture of the MMA list kernel (decode at batch 8, Llama-2-7B Q4_K stack).  usage: gpu_ncu_cols.sh TAG
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
TAG=${1:-r01}
mkdir -p gpurun_out
BS_STACK=q4_K timeout 200 python tools/bs_probe.py 8 > gpurun_out/plain.log 2>&1 && tail -1 gpurun_out/plain.log &&
BS_STACK=q4_K timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_cols -s 1 -c 1 -o gpurun_out/${TAG}_cols python tools/bs_probe.py 8 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
