#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "few_columns" 2>&1 | tail -3
BS_STACK=q4_K timeout 200 python tools/bs_probe.py 2 8 2>&1 | tail -2
timeout 200 python tools/bs_probe.py 4 8 2>&1 | tail -2
