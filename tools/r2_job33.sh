#!/bin/bash
# 2 GPUs: fused pack-and-put halo over peer memory against the NCCL halo (bit-identity tests), bench N=2 with and without it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu -k "halo_put" > gpurun_out/j33_pytest_N${NG:-2}.log 2>&1; echo "halo_put tests rc=$?"
tail -15 gpurun_out/j33_pytest_N${NG:-2}.log
for mode in put nccl; do
flag=""; [ $mode = put ] && flag="--halo-put"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NG:-2} --master-addr 127.0.0.1 --master-port 29743 bench.py --gpus ${NG:-2} --steps 20 --warmup 5 --no-cpu --no-e2e --no-parity $flag > gpurun_out/j33_bench_N${NG:-2}_$mode.json 2> gpurun_out/j33_bench_N${NG:-2}_$mode.err; echo "bench $mode rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/j33_bench_N${NG:-2}_$mode.json').read().strip().splitlines()[-1])
    print('$mode', 'lj', d['value'], d['ms_per_step'], d.get('halo'), d['kernel_ms']['halo'], 'eam', d['eam']['value'], d['eam']['ms_per_step'], d['eam']['kernel_ms']['halo'])
except Exception as e:
    print('no line', e); print(open('gpurun_out/j33_bench_N${NG:-2}_$mode.err').read()[-1500:])
PY
done
