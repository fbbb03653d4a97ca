set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_extra_formats.py -m gpu -q > gpurun_out/r03s_pytest_extra.txt 2>&1; echo pytest=$?
tail -25 gpurun_out/r03s_pytest_extra.txt
timeout 900 python -m pytest tests/test_gpu_plugin.py -m gpu -q -k "harness" > gpurun_out/r03s_pytest_harness.txt 2>&1; echo harness=$?
tail -5 gpurun_out/r03s_pytest_harness.txt
grep -c "type_a=q4_1\|type_a=q5_0\|type_a=q5_1\|type_a=iq4_nl" gpurun_out/harness_MUL_MAT.log
grep "type_a=q4_1\|type_a=q5_0\|type_a=q5_1\|type_a=iq4_nl" gpurun_out/harness_MUL_MAT.log | grep -c "OK"
grep "type_a=q4_1\|type_a=q5_0\|type_a=q5_1\|type_a=iq4_nl" gpurun_out/harness_MUL_MAT.log | grep -v "OK" | head -5
