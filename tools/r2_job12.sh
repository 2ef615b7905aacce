# compute-sanitizer, ONE tool per gpurun call (profiling recipe): $1 = memcheck | racecheck
mkdir -p gpurun_out
TOOL=${1:-memcheck}
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j12_plain.log 2>&1 || { echo "plain smoke failed"; tail -5 gpurun_out/r2j12_plain.log; exit 1; }
timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j12_${TOOL}_smoke.log 2>&1; echo "$TOOL smoke rc=$?"; tail -4 gpurun_out/r2j12_${TOOL}_smoke.log
timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 python -m pytest tests/test_zz_gpu_chain.py -m gpu -x -q -k "setup_forces or error_paths" > gpurun_out/r2j12_${TOOL}_chain.log 2>&1; echo "$TOOL chain rc=$?"; tail -6 gpurun_out/r2j12_${TOOL}_chain.log
