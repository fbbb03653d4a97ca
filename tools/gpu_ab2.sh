#!/bin/bash
# one-warp-row group count A/B (GGML_B200_BATCH_G1) on the stacks + the shape table
for g1 in 12 14; do
  echo "=== G1=$g1"; export GGML_B200_BATCH_G1=$g1
  timeout 120 python tools/batch_probe.py 20 q4_K 2>&1 | tail -1
  timeout 120 python tools/batch_probe.py 20 l3mix 2>&1 | tail -1
  timeout 120 python tools/batch_probe.py 20 q6_K 2>&1 | tail -1
  timeout 120 python tools/batch_probe.py 20 q8_0 2>&1 | tail -1
done
unset GGML_B200_BATCH_G1
timeout 300 python tools/shape_probe.py q4_K q6_K 2>&1 | tail -12
