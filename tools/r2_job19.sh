#!/bin/bash
# list build rewritten (grouped records, one pointer per group, clamp-free appends, unique-image bond path): whole GPU suite,
# kernel times at 16.4 M atoms for 128- and 256-thread blocks, the bench without its CPU legs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/j19_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/j19_pytest.log
tail -4 gpurun_out/j19_pytest.log
timeout 600 python tools/kbench.py lj 160 build128 > gpurun_out/j19_kbench_lj128.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j19_kbench_lj128.log
timeout 600 python tools/kbench.py lj 160 build256 build_threads=256 > gpurun_out/j19_kbench_lj256.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j19_kbench_lj256.log
timeout 600 python tools/kbench.py eam 100 build256 build_threads=256 > gpurun_out/j19_kbench_eam256.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|rebuild:" gpurun_out/j19_kbench_eam256.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-parity --no-cpu --no-e2e > gpurun_out/j19_bench.json 2> gpurun_out/j19_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j19_bench.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['kernel_ms']); print('eam', d['eam']['value'], d['eam']['kernel_ms']); print('chain', d['chain']['value'], d['chain']['kernel_ms'])
PY
