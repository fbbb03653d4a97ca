set -x
mkdir -p gpurun_out
T=${1:-r03g}
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.txt 2>&1; echo pytest=$?
tail -8 gpurun_out/${T}_pytest_gpu.txt
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench=$?
tail -3 gpurun_out/${T}_bench.err
