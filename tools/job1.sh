set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/j1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j1_pytest.log
for cl in 1 2 4; do B200MD_CLUSTER=$cl timeout 300 python tools/kbench.py lj 100 cl$cl >> gpurun_out/j1_kbench.log 2>&1; done
timeout 300 python tools/kbench.py eam 100 eam >> gpurun_out/j1_kbench.log 2>&1
timeout 300 python tools/kbench.py lj 80 prof > gpurun_out/j1_plain_lj.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_lj_cluster -s 70 -c 2 -o gpurun_out/j1_lj python tools/kbench.py lj 80 prof > gpurun_out/j1_ncu_lj.log 2>&1
timeout 300 python tools/kbench.py eam 60 prof > gpurun_out/j1_plain_eam.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_eam -s 140 -c 4 -o gpurun_out/j1_eam python tools/kbench.py eam 60 prof > gpurun_out/j1_ncu_eam.log 2>&1
tail -3 gpurun_out/j1_pytest.log; cat gpurun_out/j1_kbench.log
