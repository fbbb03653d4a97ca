#!/bin/bash
for lib in "" variants/ctas3.so variants/ctas4.so; do
  echo "=== lib: ${lib:-in-tree 2 CTAs/SM}"
  export GGML_B200_KERNELS_LIB=${lib:+$PWD/$lib}
  BS_STACK=q4_K timeout 200 python tools/bs_probe.py 2 8 2>&1 | tail -2
done
