# Round 2, first GPU call (one B200, about 12 box-minutes): everything written after round 1's GPU budget ran out.
#   gpurun --timeout 1500 -- 'bash tools/r2_job1.sh'
# Each step has its own timeout; logs land in gpurun_out/r2j1_*.  Read r2j1_pytest.log for x / X marks of the
# FIRST_RUN tests (tests/conftest.py): X = passed on hardware -> promote to a plain test.
mkdir -p gpurun_out; rm -f gpurun_out/r2j1_*
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > gpurun_out/r2j1_gpu.csv
timeout 900 python -m pytest tests -m gpu -q -rxX > gpurun_out/r2j1_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r2j1_pytest.log
# the opt-in variants (three-stream host pipeline, halo/compute overlap) in a process of their own
B200MD_TEST_PIPELINE=1 B200MD_TEST_OVERLAP=1 B200MD_TEST_TILED=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -rxX -k "host_buffer or overlapped or tiled" > gpurun_out/r2j1_optin.log 2>&1; echo "optin rc=$?"; tail -5 gpurun_out/r2j1_optin.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2j1_smoke.log 2>&1; echo "smoke rc=$?"
timeout 420 python bench.py --steps 100 --warmup 20 --chain 8 > gpurun_out/r2j1_bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r2j1_bench.log | cut -c1-600
# the shared-memory tiled lj/cut and eam kernels against the default ones, bins per block swept (no CPU / e2e legs)
for g in 0 4 8 12; do   # 0 = automatic (about 112 atoms per block)
  B200MD_LJ_TILED=1 B200MD_TILED_G=$g timeout 300 python bench.py --steps 100 --warmup 20 --no-cpu --no-e2e > gpurun_out/r2j1_tiled_g$g.log 2>&1
  echo "tiled G=$g rc=$?"; tail -1 gpurun_out/r2j1_tiled_g$g.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); print(l['value'], l['ms_per_step'], l.get('eam',{}).get('value'))"
done
# the e2e leg through the three-stream pipeline (ranged force kernels), against r2j1_bench.log's serial e2e
B200MD_HOST_PIPELINE=1 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --no-eam > gpurun_out/r2j1_pipeline.log 2>&1
echo "pipeline rc=$?"; tail -1 gpurun_out/r2j1_pipeline.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); print(l['value'], l['e2e'])"
timeout 120 tools/ubench > gpurun_out/r2j1_ubench.log 2>&1; echo "ubench rc=$?"; cat gpurun_out/r2j1_ubench.log
