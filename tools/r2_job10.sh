mkdir -p gpurun_out; rm -f gpurun_out/r2j10_*
timeout 600 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2j10_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j10_pytest.log
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 5 --no-parity > gpurun_out/r2j10_bench2.log 2>&1
echo "bench2 rc=$?"; tail -1 gpurun_out/r2j10_bench2.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); print(l['value'], l['ms_per_step'], l['kernel_ms']); e=l['eam']; print(e['value'], e['ms_per_step'], e['kernel_ms'])"
