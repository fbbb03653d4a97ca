# usage (inside one gpurun call, 1 GPU): bash tools/job_sanitizer.sh TAG
# compute-sanitizer memcheck over one parametrization of each kernel family (the sanitizer runs kernels 10-100x slower:
# every leg has its own timeout and the selection is small)
set -x
T=${1:-rXX}
mkdir -p gpurun_out
S="compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20"
timeout 600 $S python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "few_columns and 2 and not long" > gpurun_out/${T}_memcheck_cols.txt 2>&1; echo cols=$?
timeout 300 $S python -m pytest tests/test_gpu_extra_formats.py -m gpu -q -x -k "quantize_rows" > gpurun_out/${T}_memcheck_quantize_rows.txt 2>&1; echo qrows=$?
timeout 600 $S python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_memcheck_smoke.txt 2>&1; echo smoke=$?
tail -3 gpurun_out/${T}_memcheck_*.txt
