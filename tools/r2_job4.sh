# Round 2 GPU call: ncu --set full of the tiled lj kernel next to the default one (2 M atoms), plain run first.
#   gpurun --timeout 1500 -- 'bash tools/r2_job4.sh'
mkdir -p gpurun_out; rm -f gpurun_out/r2j4_*
for mode in tiled default; do
  if [ $mode = tiled ]; then export B200MD_LJ_TILED=1; else unset B200MD_LJ_TILED; fi
  timeout 300 python tools/kbench.py lj 80 $mode > gpurun_out/r2j4_plain_$mode.log 2>&1 || { echo "plain $mode failed"; tail -5 gpurun_out/r2j4_plain_$mode.log; continue; }
  tail -4 gpurun_out/r2j4_plain_$mode.log
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_lj_tiled|k_lj_lanes|k_build_bins" \
      -c 5 -s 10 -o gpurun_out/r2j4_lj_$mode python tools/kbench.py lj 80 $mode > gpurun_out/r2j4_ncu_$mode.log 2>&1
  echo "ncu $mode rc=$?"
done
