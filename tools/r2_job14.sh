# Round 2 GPU call (1 GPU): full pytest, the default bench line, the ncu launch list of the timed command, ncu of the chain build
mkdir -p gpurun_out; rm -f gpurun_out/r2j14_*
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2j14_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j14_pytest.log
timeout 1200 python bench.py --steps 20 --warmup 5 > gpurun_out/r2j14_bench.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r2j14_bench.log | cut -c1-200
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-e2e --no-parity"
timeout 400 $CMD > gpurun_out/r2j14_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2j14_launches.csv $CMD > gpurun_out/r2j14_ncu_list.log 2>&1
echo "launch list rc=$?"
CMD2="python bench.py --steps 5 --warmup 3 --cells 80 --no-eam --chain 4 --no-cpu --no-e2e --no-parity"
timeout 300 $CMD2 > gpurun_out/r2j14_plain2.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_build_seg" -s 1 -c 2 -o gpurun_out/r2j14_chainbuild $CMD2 > gpurun_out/r2j14_ncu2.log 2>&1
echo "ncu chain build rc=$?"
