#!/bin/bash
# final: launch list of the bench command (kernel shares), then the full default bench line
mkdir -p gpurun_out; rm -f gpurun_out/j29_*
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-e2e --no-parity"
timeout 600 $CMD > gpurun_out/j29_plain.json 2> gpurun_out/j29_plain.err || { echo "plain run failed"; tail -5 gpurun_out/j29_plain.err; exit 1; }
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/j29_launches.csv $CMD > gpurun_out/j29_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/j29_launches.csv
timeout 1500 python bench.py > gpurun_out/j29_bench_default.json 2> gpurun_out/j29_bench_default.err; echo "default bench (no flags) rc=$?"
tail -c 300 gpurun_out/j29_bench_default.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/j29_bench_reference.json 2> gpurun_out/j29_bench_reference.err; echo "reference arm rc=$?"
tail -c 600 gpurun_out/j29_bench_reference.json
