#!/bin/bash
# copy-back in the host's order on the device (b200md_download_atoms_by_tag): whole GPU suite, then the bench with its e2e legs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/j30_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/j30_pytest.log
tail -4 gpurun_out/j30_pytest.log
timeout 1500 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/j30_bench.json 2> gpurun_out/j30_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j30_bench.json').read().strip().splitlines()[-1])
print('lj', d['value'], 'e2e', d['e2e']['value'], 'lammps_host', d['e2e'].get('lammps_host'), 'parity', d.get('parity_all_ok'))
PY
