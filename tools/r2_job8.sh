# Round 2 multi-GPU call (N = 2): decomposition tests (overlap is the default on several ranks), bench with the parity legs
mkdir -p gpurun_out; rm -f gpurun_out/r2j8_*
nvidia-smi -L | head -3
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/r2j8_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j8_pytest.log
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2j8_bench2.log 2>&1
echo "bench2 rc=$?"; tail -1 gpurun_out/r2j8_bench2.log | cut -c1-400
