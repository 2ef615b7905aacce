#!/bin/bash
# fix nve fused into the force kernels + parallel build prologue: whole GPU suite, kernel times, bench without its CPU legs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/j21_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/j21_pytest.log
tail -4 gpurun_out/j21_pytest.log
timeout 600 python tools/kbench.py lj 160 fused > gpurun_out/j21_kbench_lj.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j21_kbench_lj.log
timeout 600 python tools/kbench.py lj 160 unfused fuse_nve=0 > gpurun_out/j21_kbench_lj0.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j21_kbench_lj0.log
timeout 600 python tools/kbench.py eam 100 fused > gpurun_out/j21_kbench_eam.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j21_kbench_eam.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-parity --no-cpu --no-e2e > gpurun_out/j21_bench.json 2> gpurun_out/j21_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j21_bench.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['ms_per_step'], d['kernel_ms']); print('eam', d['eam']['value'], d['eam']['kernel_ms']); print('chain', d['chain']['value'], d['chain']['kernel_ms'])
PY
