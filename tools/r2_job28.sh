#!/bin/bash
# weak-scaling line at N GPUs (N = $1) with the fused kernels: multirank tests that fit, then bench --gpus N
N=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu > gpurun_out/j28_pytest_N$N.log 2>&1; echo "multirank tests rc=$?"
tail -3 gpurun_out/j28_pytest_N$N.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29733 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu > gpurun_out/j28_bench_N$N.json 2> gpurun_out/j28_bench_N$N.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/j28_bench_N$N.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['ms_per_step'], d['kernel_ms'], d.get('parity',{}).get('ok')); print('eam', d['eam']['value'], d['eam']['ms_per_step'], d['eam'].get('parity',{}).get('ok')); print('parity_all_ok', d.get('parity_all_ok'), 'e2e', (d.get('e2e') or {}).get('value'))
PY
