set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_peers.py -m gpu -q -k "two_devices or peers or fused or split" > gpurun_out/r03zz_pytest_2gpu.txt 2>&1; echo pytest=$?
tail -8 gpurun_out/r03zz_pytest_2gpu.txt
bash tools/job_bench_ngpu.sh 2 2>&1 | tail -8
