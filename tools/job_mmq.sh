set -x
mkdir -p gpurun_out
T=${1:-r03j}
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mul_mat_q or prefill or configs or tensor or scale or route or int8 or smoke" > gpurun_out/${T}_pytest_mmq.txt 2>&1; echo pytest=$?
tail -6 gpurun_out/${T}_pytest_mmq.txt
timeout 600 python tools/prefill_l3_probe.py > gpurun_out/${T}_prefill_l3.txt 2>&1; cat gpurun_out/${T}_prefill_l3.txt | tail -20
GGML_B200_MMQ_BN=256 timeout 600 python tools/prefill_l3_probe.py 2>&1 | tail -20 > gpurun_out/${T}_prefill_l3_bn256.txt; cat gpurun_out/${T}_prefill_l3_bn256.txt
timeout 600 python tools/prefill_probe.py 2>&1 | grep "n=  512\|n= 2048" > gpurun_out/${T}_prefill.txt; cat gpurun_out/${T}_prefill.txt
