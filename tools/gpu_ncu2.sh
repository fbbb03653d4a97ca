#!/bin/bash
# ncu evidence for the bench command: launch list of our kernels + one full capture.  usage: gpu_ncu2.sh TAG
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
TAG=${1:-r01}
mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/plain2.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_mmvq|k_quantize|k_mmq|k_deq' -c 60 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/ncu2.log 2>&1
tail -4 gpurun_out/${TAG}_launches.csv | cut -c1-300
timeout 120 python tools/batch_probe.py 0 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_batch -s 1 -c 1 -o gpurun_out/${TAG}_batch python tools/batch_probe.py 0 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
