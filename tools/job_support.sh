set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "dequant or quantize or get_rows or cpy or golden" > gpurun_out/r03n_pytest_support.txt 2>&1; echo pytest=$?
tail -4 gpurun_out/r03n_pytest_support.txt
timeout 600 python tools/support_probe.py > gpurun_out/r03n_support_probe.txt 2>&1; cat gpurun_out/r03n_support_probe.txt
