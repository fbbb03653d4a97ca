#!/bin/bash
# round-1 status check on the GPU box: tests, probes, bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu.log 2>&1
( timeout 200 tools/stream_probe ) > gpurun_out/stream_probe.log 2>&1
for shape in "4096 4096" "11008 4096" "4096 11008" "32000 4096"; do
  for nf in 0 1; do
    echo -n "no_fused=$nf " >> gpurun_out/mmvq_probe.log
    GGML_B200_MMVQ_NO_FUSED=$nf timeout 120 python tools/mmvq_probe.py q4_K $shape >> gpurun_out/mmvq_probe.log 2>&1
  done
done
( timeout 600 python bench.py --steps 30 --warmup 3 ) > gpurun_out/bench1.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/mmvq_probe.log; tail -2 gpurun_out/bench1.log
