set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "columns or few_columns or batch" > gpurun_out/r03e_pytest_cols.txt 2>&1; echo pytest=$?
tail -5 gpurun_out/r03e_pytest_cols.txt
BS_STACK=q4_K timeout 300 python tools/bs_probe.py 2 8 2>&1 | tail -2
BS_STACK=q5_K timeout 300 python tools/bs_probe.py 8 2>&1 | tail -1
timeout 300 python tools/bs_probe.py 8 2>&1 | tail -1
