# Round 2 GPU call: the thread-per-atom / staged-candidate list build (k_build_seg): tests, timings at 16.4 M, ncu
mkdir -p gpurun_out; rm -f gpurun_out/r2j7_*
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2j7_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2j7_pytest.log
timeout 300 python tools/kbench.py lj 160 seg > gpurun_out/r2j7_kbench_lj.log 2>&1; tail -4 gpurun_out/r2j7_kbench_lj.log
timeout 300 python tools/kbench.py eam 100 seg > gpurun_out/r2j7_kbench_eam.log 2>&1; tail -4 gpurun_out/r2j7_kbench_eam.log
CMD="python bench.py --steps 20 --warmup 5 --cells 80 --no-eam --chain 8 --no-cpu --no-e2e --no-parity"
timeout 300 $CMD > gpurun_out/r2j7_chain.log 2>&1 && tail -1 gpurun_out/r2j7_chain.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); c=l['chain']; print('chain', c['value'], c['ms_per_step'], c['kernel_ms'])"
CMD2="python bench.py --steps 5 --warmup 3 --cells 80 --no-eam --chain 4 --no-cpu --no-e2e --no-parity"
timeout 300 $CMD2 > gpurun_out/r2j7_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_build_seg" -c 4 -o gpurun_out/r2j7_build $CMD2 > gpurun_out/r2j7_ncu.log 2>&1
echo "ncu rc=$?"
