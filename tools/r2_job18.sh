#!/bin/bash
# EAM with the in-range lists (rho pass leaves the entries inside the force cutoff for the force pass): parity tests, kernel times
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -k "eam" > gpurun_out/j18_pytest_eam.log 2>&1; echo "eam tests rc=$?" | tee -a gpurun_out/j18_pytest_eam.log
tail -4 gpurun_out/j18_pytest_eam.log
timeout 600 python tools/kbench.py eam 100 inrange > gpurun_out/j18_kbench_eam.log 2>&1; echo "kbench rc=$?"
grep -a "kbench\|pair:\|list:" gpurun_out/j18_kbench_eam.log
timeout 600 python tools/kbench.py eam 100 inrange_overlap overlap=1 > gpurun_out/j18_kbench_eam_ov.log 2>&1; echo "kbench ov rc=$?"
grep -a "kbench\|pair:" gpurun_out/j18_kbench_eam_ov.log
