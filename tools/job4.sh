mkdir -p gpurun_out; rm -f gpurun_out/j4_kbench.log
for cl in 1 2 4; do B200MD_LANES=1 B200MD_CLUSTER=$cl timeout 300 python tools/kbench.py lj 100 lanes$cl >> gpurun_out/j4_kbench.log 2>&1; done
B200MD_LANES=1 B200MD_CLUSTER=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lj" >> gpurun_out/j4_kbench.log 2>&1
grep "kbench\|passed\|failed" gpurun_out/j4_kbench.log
