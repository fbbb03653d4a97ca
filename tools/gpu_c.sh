#!/bin/bash


timeout 200 python tools/bs_probe.py 2>&1 | tail -5
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "persistent or batch" 2>&1 | tail -4
