set -x
mkdir -p gpurun_out
BS_STACK=q4_K timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_cols_ring -s 2 -c 1 -o gpurun_out/r03i_ring_q4k python tools/bs_probe.py 8 > gpurun_out/r03i_ncu.log 2>&1; echo ncu=$?
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmvq_cols_ringILi1 -s 2 -c 1 -o gpurun_out/r03i_ring_q6k python tools/bs_probe.py 8 > gpurun_out/r03i_ncu6.log 2>&1; echo ncu6=$?
tail -3 gpurun_out/r03i_ncu6.log
