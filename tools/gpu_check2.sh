#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "persistent or batch or fused" 2>&1 | tail -15 ) > gpurun_out/pytest_batch.log 2>&1
tail -5 gpurun_out/pytest_batch.log
for cfg in "9216 16 3" "9216 8 3" "18432 8 3" "9216 16 2" "4608 16 3"; do set -- $cfg
  echo -n "slot=$1 stages=$2 abuf=$3: "
  GGML_B200_BATCH_SLOT_BYTES=$1 GGML_B200_BATCH_STAGES=$2 GGML_B200_BATCH_ABUF=$3 timeout 120 python tools/batch_probe.py 20 2>&1 | tail -1
done | tee gpurun_out/batch_probe.log
