#!/bin/bash
# usage: gpu_multi.sh N [steps]
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}; K=${2:-100}
BENCH_DEBUG=1 NCCL_DEBUG=WARN timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps $K --warmup 5 $BENCH_ARGS > gpurun_out/multi_$N.json 2> gpurun_out/multi_$N.err
echo rc=$?; grep "bench r" gpurun_out/multi_$N.err | tail -20; cat gpurun_out/multi_$N.json | cut -c1-1500
