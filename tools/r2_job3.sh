# Round 2, multi-GPU call (charged N x box time: keep it short).  The e2e leg that hung round 1's 4- and 8-GPU runs is
# bounded by bench.py's own watchdog now; the outer timeouts are the second line of defence.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/r2_job3.sh'
mkdir -p gpurun_out; rm -f gpurun_out/r2j3_*
timeout 420 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/r2j3_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j3_pytest.log
for n in 8 4 2; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 100 --warmup 20 --no-cpu > gpurun_out/r2j3_bench$n.log 2>&1
  echo "bench$n rc=$?"; tail -1 gpurun_out/r2j3_bench$n.log | cut -c1-500
done
