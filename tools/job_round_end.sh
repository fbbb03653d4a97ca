# usage (inside one gpurun call, 1 GPU): bash tools/job_round_end.sh TAG
# the round's evidence set: GPU tests, smoke, bench (ours + reference arm), ncu launch list of the bench step, full ncu
# captures of the dominant kernel (k_mmvq_batch) and of the 2..8-column kernel; probes of the secondary figures
set -x
T=${1:-rXX}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${T}_gpu.txt
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.txt 2>&1; echo pytest=$?
tail -4 gpurun_out/${T}_pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.txt 2>&1; echo smoke=$?
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo bench=$?
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; echo ref=$?
timeout 300 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/${T}_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/${T}_ncu1.log 2>&1; echo launches=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_batch -s 8 -c 1 -o gpurun_out/${T}_batch python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/${T}_ncu2.log 2>&1; echo full=$?
timeout 300 python tools/bs_probe.py 1 2 4 8 > gpurun_out/${T}_bs_probe_l3mix.txt 2>&1
BS_STACK=q4_K timeout 300 python tools/bs_probe.py 2 4 8 > gpurun_out/${T}_bs_probe_q4k.txt 2>&1
for S in q5_K q6_K q4_0 q8_0; do BS_STACK=$S timeout 300 python tools/bs_probe.py 1 2 8 2>&1 | tail -3; done > gpurun_out/${T}_bs_probe_other_formats.txt
timeout 600 python tools/prefill_l3_probe.py > gpurun_out/${T}_prefill_l3.txt 2>&1
timeout 600 python tools/support_probe.py > gpurun_out/${T}_support_probe.txt 2>&1
