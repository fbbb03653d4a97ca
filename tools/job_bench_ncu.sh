set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02v_gpu.txt
timeout 600 python bench.py > gpurun_out/r02v_bench.json 2> gpurun_out/r02v_bench.err; echo bench=$?
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02v_bench_reference.json 2> gpurun_out/r02v_bench_reference.err; echo ref=$?
timeout 300 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/r02v_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file gpurun_out/r02v_launches.csv python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/r02v_ncu1.log 2>&1; echo launches=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_batch -s 8 -c 1 -o gpurun_out/r02v_batch python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/r02v_ncu2.log 2>&1; echo full=$?
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02v_smoke.txt 2>&1; echo smoke=$?
