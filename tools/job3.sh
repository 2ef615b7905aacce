set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/j3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j3_pytest.log
tail -15 gpurun_out/j3_pytest.log
timeout 300 python tools/kbench.py lj 100 new > gpurun_out/j3_kbench.log 2>&1
B200MD_BUILD_FP64=1 timeout 300 python tools/kbench.py lj 100 oldbuild >> gpurun_out/j3_kbench.log 2>&1
timeout 300 python tools/kbench.py eam 100 new >> gpurun_out/j3_kbench.log 2>&1
cat gpurun_out/j3_kbench.log
