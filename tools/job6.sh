mkdir -p gpurun_out; rm -f gpurun_out/j6_*.log
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/j6_pytest.log 2>&1; tail -5 gpurun_out/j6_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 20 --no-cpu > gpurun_out/j6_bench2.log 2>&1; tail -2 gpurun_out/j6_bench2.log | cut -c1-3000
