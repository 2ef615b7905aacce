#!/bin/bash
# ncu --set full of the rewritten list build (LJ melt 2 M atoms, chain 2 M beads) and of the EAM passes with the in-range lists
mkdir -p gpurun_out; rm -f gpurun_out/j20_*
CMD="python bench.py --steps 5 --warmup 3 --cells 80 --eam-cells 60 --chain 4 --no-cpu --no-e2e --no-parity"
timeout 600 python -m pytest tests/test_lammps_host.py -x -q -m gpu -k "device_list" > gpurun_out/j20_listtest.log 2>&1; tail -3 gpurun_out/j20_listtest.log
timeout 300 $CMD > gpurun_out/j20_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/j20_plain.log; exit 1; }
tail -1 gpurun_out/j20_plain.log | cut -c1-200
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"k_build_seg|k_eam_rho_fp_1t|k_eam_force_1t" -c 12 -o gpurun_out/j20_build_eam $CMD > gpurun_out/j20_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/j20_ncu.log
ls -la gpurun_out/
