#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "few_columns" 2>&1 | tail -3
GGML_B200_COLS_KS=1 timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "few_columns" 2>&1 | tail -3
for ks in 8 1; do echo "== KS=$ks"; GGML_B200_COLS_KS=$ks BS_STACK=q4_K timeout 200 python tools/bs_probe.py 2 8 2>&1 | tail -2; done
timeout 200 python tools/bs_probe.py 4 8 2>&1 | tail -2
