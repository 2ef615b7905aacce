# 1 GPU: does the overlap split itself cost time (periodic self-halo, no NCCL)?  + ncu of the split launches
mkdir -p gpurun_out; rm -f gpurun_out/r2j11_*
timeout 300 python tools/kbench.py lj 160 plain > gpurun_out/r2j11_plain.log 2>&1; tail -4 gpurun_out/r2j11_plain.log | head -1
timeout 300 python tools/kbench.py lj 160 overlap overlap=1 > gpurun_out/r2j11_overlap.log 2>&1; tail -4 gpurun_out/r2j11_overlap.log | head -1
timeout 300 python tools/kbench.py eam 100 overlap overlap=1 > gpurun_out/r2j11_overlap_eam.log 2>&1; tail -4 gpurun_out/r2j11_overlap_eam.log | head -1
timeout 300 python tools/kbench.py lj 80 overlap overlap=1 > gpurun_out/r2j11_ov80.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"k_lj_force" -s 20 -c 4 -o gpurun_out/r2j11_split python tools/kbench.py lj 80 overlap overlap=1 > gpurun_out/r2j11_ncu.log 2>&1
echo "ncu rc=$?"
