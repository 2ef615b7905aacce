#!/bin/bash
# 2 GPUs: decomposition tests (NCCL halo, migration, overlap default, fused nve in the split force launches), bench at N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu > gpurun_out/j23_pytest.log 2>&1; echo "multirank tests rc=$?" | tee -a gpurun_out/j23_pytest.log
tail -4 gpurun_out/j23_pytest.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu > gpurun_out/j23_bench_N2.json 2> gpurun_out/j23_bench_N2.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/j23_bench_N2.json').read().strip().splitlines()[-1])
print('lj', d['value'], d['ms_per_step'], d['kernel_ms'], d.get('parity',{}).get('ok')); print('eam', d['eam']['value'], d['eam']['ms_per_step'], d['eam'].get('parity',{}).get('ok')); print('parity_all_ok', d.get('parity_all_ok'))
PY
