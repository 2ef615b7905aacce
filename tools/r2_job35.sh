#!/bin/bash
# rebuild-side fusions (bin padding written by k_rec_fill, xhold by the permutation kernel): whole GPU suite + bench without CPU legs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/j35_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/j35_pytest.log
tail -3 gpurun_out/j35_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-parity --no-cpu --no-e2e > gpurun_out/j35_bench.json 2> gpurun_out/j35_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j35_bench.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['ms_per_step']); print('eam', d['eam']['value'], d['eam']['ms_per_step']); print('chain', d['chain']['value'], d['chain']['ms_per_step'], d['chain']['kernel_ms'])
PY
