#!/bin/bash
# list build with two atoms per thread: whole GPU suite, kernel times for 1 and 2 atoms per thread, bench without CPU legs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/j24_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/j24_pytest.log
tail -4 gpurun_out/j24_pytest.log
for na in 2 1; do
timeout 600 python tools/kbench.py lj 160 na$na build_na=$na > gpurun_out/j24_kbench_lj_na$na.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j24_kbench_lj_na$na.log
timeout 600 python tools/kbench.py eam 100 na$na build_na=$na > gpurun_out/j24_kbench_eam_na$na.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j24_kbench_eam_na$na.log
done
timeout 600 python tools/kbench.py lj 160 na2_256 build_na=2,build_threads=256 > gpurun_out/j24_kbench_lj_na2_256.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j24_kbench_lj_na2_256.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-parity --no-cpu --no-e2e > gpurun_out/j24_bench.json 2> gpurun_out/j24_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j24_bench.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['ms_per_step'], d['kernel_ms'], d['roofline']['frac'], d['roofline'].get('rows')); print('eam', d['eam']['value'], d['eam']['kernel_ms']); print('chain', d['chain']['value'], d['chain']['kernel_ms'])
PY
