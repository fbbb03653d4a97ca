#!/bin/bash
# usage: gpu_check2.sh ["slot stages abuf" ...]
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "persistent or batch or fused" 2>&1 | tail -15 ) > gpurun_out/pytest_batch.log 2>&1
tail -3 gpurun_out/pytest_batch.log
[ $# -eq 0 ] && set -- "36864 16 3" "46080 16 2"
for cfg in "$@"; do set -- $cfg
  echo -n "slot=$1 stages=$2 abuf=$3: "
  GGML_B200_BATCH_SLOT_BYTES=$1 GGML_B200_BATCH_STAGES=$2 GGML_B200_BATCH_ABUF=$3 timeout 120 python tools/batch_probe.py 20 2>&1 | tail -1
done | tee gpurun_out/batch_probe.log
