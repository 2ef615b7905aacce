# Round 2 GPU call: the refactored defaults (host pipeline, no tiled kernels), promoted tests, bench with parity legs.
mkdir -p gpurun_out; rm -f gpurun_out/r2j5_*
timeout 900 python -m pytest tests -m gpu -q -x -rxXs > gpurun_out/r2j5_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2j5_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j5_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2j5_bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r2j5_bench.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2j5_bench_ref.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/r2j5_bench_ref.log | cut -c1-600
