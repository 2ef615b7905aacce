#!/bin/bash
# in.chain: bonds + thermostat + fix nve as one pass (k_chain_post): chain tests on the GPU, bench without its CPU legs
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_zz_gpu_chain.py tests/test_lammps_host.py -x -q -m gpu -k "chain" > gpurun_out/j22_pytest.log 2>&1; echo "chain tests rc=$?" | tee -a gpurun_out/j22_pytest.log
tail -4 gpurun_out/j22_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-parity --no-cpu --no-e2e --no-eam > gpurun_out/j22_bench.json 2> gpurun_out/j22_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j22_bench.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['ms_per_step'], d['kernel_ms']); print('chain', d['chain']['value'], d['chain']['ms_per_step'], d['chain']['kernel_ms'])
PY
