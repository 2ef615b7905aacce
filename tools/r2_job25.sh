#!/bin/bash
# final captures of the kernels that run: launch list of the bench command, ncu --set full of each hot kernel, then the full default bench
mkdir -p gpurun_out; rm -f gpurun_out/j25_*
CMD="python bench.py --steps 5 --warmup 3 --cells 80 --eam-cells 60 --chain 4 --no-cpu --no-e2e --no-parity"
timeout 300 $CMD > gpurun_out/j25_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/j25_plain.log; exit 1; }
tail -1 gpurun_out/j25_plain.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/j25_launches.csv $CMD > gpurun_out/j25_launches.log 2>&1
echo "launch list rc=$?"
for k in k_lj_force k_eam_rho_fp_1t k_eam_force_1t k_build_seg k_chain_post k_bond_fene; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$k" -s 6 -c 3 -o gpurun_out/j25_$k $CMD > gpurun_out/j25_ncu_$k.log 2>&1
echo "ncu $k rc=$?"
done
ls -la gpurun_out/ | grep j25
timeout 1500 python bench.py --steps 20 --warmup 5 > gpurun_out/j25_bench_full.json 2> gpurun_out/j25_bench_full.err; echo "full bench rc=$?"
tail -c 600 gpurun_out/j25_bench_full.json
