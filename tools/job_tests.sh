set -x
mkdir -p gpurun_out
T=${1:-r03h}
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.txt 2>&1; echo pytest=$?
tail -15 gpurun_out/${T}_pytest_gpu.txt
