# Round 2 multi-GPU call (N = 8, then 4): decomposition tests + bench with the parity legs (outer timeouts bound a hang)
mkdir -p gpurun_out; rm -f gpurun_out/r2j13_*
nvidia-smi -L | wc -l
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/r2j13_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j13_pytest.log
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2j13_bench$n.log 2>&1
  echo "bench$n rc=$?"; tail -1 gpurun_out/r2j13_bench$n.log | python -c "
import sys,json
l=json.loads(sys.stdin.read()); e=l['eam']
print(l['n_gpus'], l['value'], l['ms_per_step'], l['kernel_ms'], 'e2e', l['e2e']['value'] if l.get('e2e') else None, 'parity', l.get('parity_all_ok'))
print('eam', e['value'], e['ms_per_step'], e['kernel_ms'])"
done
