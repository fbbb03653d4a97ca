set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gguf.py tests/test_gpu_plugin.py -k "gguf" -m gpu -x -q > gpurun_out/r03a_pytest_gguf.txt 2>&1; echo pytest=$?
tail -15 gpurun_out/r03a_pytest_gguf.txt
timeout 300 python - > gpurun_out/r03a_gguf_load.txt 2>&1 <<'PY'
import sys, json
sys.path.insert(0, "ggml-gfx906_b200"); sys.path.insert(0, "tests")
import torch, b200, bench
kern = b200.Kernels()
print(json.dumps(bench.measure_gguf_load(torch, kern, gib=2.0)))
PY
cat gpurun_out/r03a_gguf_load.txt
