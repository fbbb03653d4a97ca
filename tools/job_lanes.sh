set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_plugin.py -m gpu -x -q -k "lanes or batch or decoder_layer or decode_graph or graph_mul_mat" > gpurun_out/r03t_pytest_lanes.txt 2>&1; echo pytest=$?
tail -12 gpurun_out/r03t_pytest_lanes.txt
timeout 600 python tools/prefill_l3_probe.py 2>&1 | grep "one layer"
GGML_B200_LIST_LANES=0 timeout 600 python tools/prefill_l3_probe.py 2>&1 | grep "one layer"
