set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "columns or few_columns or batch" > gpurun_out/r03b_pytest_cols.txt 2>&1; echo pytest=$?
tail -30 gpurun_out/r03b_pytest_cols.txt
BS_STACK=q4_K timeout 300 python tools/bs_probe.py 2 8 > gpurun_out/r03b_bs_probe_q4k.txt 2>&1; cat gpurun_out/r03b_bs_probe_q4k.txt | tail -4
BS_STACK=q5_K timeout 300 python tools/bs_probe.py 8 > gpurun_out/r03b_bs_probe_q5k.txt 2>&1; cat gpurun_out/r03b_bs_probe_q5k.txt | tail -4
timeout 300 python tools/bs_probe.py 1 2 8 > gpurun_out/r03b_bs_probe_l3mix.txt 2>&1; cat gpurun_out/r03b_bs_probe_l3mix.txt | tail -4
