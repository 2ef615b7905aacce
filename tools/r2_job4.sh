# Round 2, fourth GPU call (one B200): ncu --set full of the tiled pair kernels next to the default ones, only after
# r2_job1 has shown the tiled path green and the same commands exit 0 without ncu.
#   gpurun --timeout 1800 -- 'bash tools/r2_job4.sh'
# Read here with:  ncu -i gpurun_out/r2j4_lj.ncu-rep --page raw --csv | python tools/ncu_summary.py
mkdir -p gpurun_out; rm -f gpurun_out/r2j4_*
export B200MD_LJ_TILED=1
for kind in "lj 80" "eam 60"; do
  tagk=${kind%% *}
  timeout 300 python tools/kbench.py $kind tiled > gpurun_out/r2j4_plain_$tagk.log 2>&1 || { echo "plain $tagk failed"; tail -5 gpurun_out/r2j4_plain_$tagk.log; continue; }
  tail -3 gpurun_out/r2j4_plain_$tagk.log
  timeout 700 ncu --set full --clock-control none --import-source on -k regex:"k_lj_tiled|k_eam_rho_fp_tiled|k_eam_force_tiled|k_build_bins" \
      -c 8 -s 12 -o gpurun_out/r2j4_$tagk python tools/kbench.py $kind tiled > gpurun_out/r2j4_ncu_$tagk.log 2>&1
  echo "ncu $tagk rc=$?"
done
