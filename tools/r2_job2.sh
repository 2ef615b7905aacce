# Round 2, second GPU call (one B200): ncu evidence for the in.chain kernels and the opt-in build staging, only after
# r2_job1 has shown the same commands exit 0 without ncu.
#   gpurun --timeout 1500 -- 'bash tools/r2_job2.sh'
mkdir -p gpurun_out; rm -f gpurun_out/r2j2_*
CMD="python bench.py --steps 20 --warmup 5 --no-eam --no-cpu --no-e2e --cells 100 --chain 6"
timeout 300 $CMD > gpurun_out/r2j2_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2j2_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2j2_launches.csv $CMD > gpurun_out/r2j2_ncu_list.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_build_bins|k_bond_fene|k_langevin_counter|k_lj_lanes" -c 12 -s 200 -o gpurun_out/r2j2_chain $CMD > gpurun_out/r2j2_ncu_full.log 2>&1; echo "full capture rc=$?"
B200MD_BUILD_STAGE=1 timeout 300 python bench.py --steps 100 --warmup 20 --no-eam --no-cpu --no-e2e > gpurun_out/r2j2_stage.log 2>&1; echo "stage rc=$?"; tail -1 gpurun_out/r2j2_stage.log | cut -c1-400
