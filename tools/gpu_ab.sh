#!/bin/bash
# A/B of a tuning build of the kernel library (variants/*.so) against the in-tree one: stack + per-shape probes.
V=${1:-variants/nwc14.so}
for lib in "" "$V"; do
  echo "=== lib: ${lib:-in-tree}"
  export GGML_B200_KERNELS_LIB=${lib:+$PWD/$lib}
  timeout 120 python tools/batch_probe.py 20 q4_K 2>&1 | tail -1
  timeout 120 python tools/batch_probe.py 20 l3mix 2>&1 | tail -1
  timeout 200 python tools/shape_probe.py q4_K ${AB_TYPES} 2>&1 | tail -12
done
