# usage: bash tools/job_bench_ngpu.sh N   (inside a gpurun --gpus N call)
N=$1
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 300 --warmup 5 > gpurun_out/r03zz_bench_${N}gpu.out 2> gpurun_out/r03zz_bench_${N}gpu.err; echo rc=$?
grep "^{" gpurun_out/r03zz_bench_${N}gpu.out > gpurun_out/r03zz_bench_${N}gpu.json
python - <<PY
import json
d=json.loads(open("gpurun_out/r03zz_bench_${N}gpu.json").read())
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"])
print(json.dumps(d.get("multi_gpu")))
print(json.dumps(d.get("decode_chain")))
PY
grep -v "^W1105\|^\*\*\*\|OMP_NUM" gpurun_out/r03zz_bench_${N}gpu.err | tail -5
