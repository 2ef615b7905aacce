mkdir -p gpurun_out; rm -f gpurun_out/j8_*
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/j8_pytest.log 2>&1; tail -3 gpurun_out/j8_pytest.log
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 --warmup 20 --no-cpu > gpurun_out/j8_bench$n.log 2>&1; echo "bench$n rc=$?"; tail -1 gpurun_out/j8_bench$n.log | cut -c1-400
done
