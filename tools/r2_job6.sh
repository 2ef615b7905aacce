# Round 2 GPU call: ncu --set full of the neighbor-build kernel on the LJ melt (2 M atoms) and the chain melt (2 M beads)
mkdir -p gpurun_out; rm -f gpurun_out/r2j6_*
CMD="python bench.py --steps 5 --warmup 3 --cells 80 --no-eam --chain 4 --no-cpu --no-e2e --no-parity"
timeout 300 $CMD > gpurun_out/r2j6_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/r2j6_plain.log; exit 1; }
tail -1 gpurun_out/r2j6_plain.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_build_bins|k_binrec_fill|k_bin_sortseg|k_bin_place|k_bin_count|k_permute_owned" -c 24 -o gpurun_out/r2j6_build $CMD > gpurun_out/r2j6_ncu.log 2>&1
echo "ncu rc=$?"
