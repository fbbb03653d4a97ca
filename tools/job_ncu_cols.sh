# usage (inside one gpurun call, 1 GPU): bash tools/job_ncu_cols.sh TAG
# full ncu captures of the 2..8-column ring kernel, one per format family, on the Llama-2-7B stack of that format at 8 columns
# (each only after the same command has exited 0 without ncu)
set -x
T=${1:-rXX}
mkdir -p gpurun_out
for S in q4_0 q6_K q4_K q8_0; do
  BS_STACK=$S timeout 300 python tools/bs_probe.py 8 > gpurun_out/${T}_cols_${S}_plain.txt 2>&1 && \
  BS_STACK=$S timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_cols_ring -s 2 -c 1 -o gpurun_out/${T}_cols_${S} python tools/bs_probe.py 8 > gpurun_out/${T}_cols_${S}_ncu.log 2>&1; echo $S=$?
done
