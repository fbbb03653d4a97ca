#!/bin/bash
# BASELINE configs[4] on the GPUs of one box: the Llama-3-70B FFN stack row-split over N ranks (strong scaling),
# plus the headline stack at the largest N.  usage: gpu_scale70.sh "2 4 8"
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for N in $1; do
  NCCL_DEBUG=WARN timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + N)) \
      bench.py --gpus $N --steps 100 --warmup 5 --stack llama3-70b-ffn --no-extras > gpurun_out/scale70_$N.json 2> gpurun_out/scale70_$N.err
  echo "N=$N rc=$?"; cut -c1-420 gpurun_out/scale70_$N.json
  last=$N
done
if [ "$last" -gt 4 ]; then
  NCCL_DEBUG=WARN timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $last --master-addr 127.0.0.1 --master-port 29540 \
      bench.py --gpus $last --steps 100 --warmup 5 > gpurun_out/multi_$last.json 2> gpurun_out/multi_$last.err
  echo "7b N=$last rc=$?"; cut -c1-420 gpurun_out/multi_$last.json
fi
