mkdir -p gpurun_out; rm -f gpurun_out/j5_*.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/j5_pytest.log 2>&1; tail -3 gpurun_out/j5_pytest.log
timeout 300 python tools/kbench.py lj 100 pipelined >> gpurun_out/j5_kbench.log 2>&1
timeout 300 python tools/kbench.py eam 100 pipelined >> gpurun_out/j5_kbench.log 2>&1
grep "kbench\|rebuild" gpurun_out/j5_kbench.log
