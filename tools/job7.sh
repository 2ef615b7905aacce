mkdir -p gpurun_out; rm -f gpurun_out/j7_*
timeout 900 python bench.py > gpurun_out/j7_bench1.json 2> gpurun_out/j7_bench1.err; echo "bench rc=$?"
tail -c 600 gpurun_out/j7_bench1.err
timeout 300 python bench.py --cells 100 --eam-cells 60 --steps 22 --warmup 3 --no-cpu --no-e2e > gpurun_out/j7_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/j7_launches.csv python bench.py --cells 100 --eam-cells 60 --steps 22 --warmup 3 --no-cpu --no-e2e > gpurun_out/j7_ncu.log 2>&1
echo "ncu rc=$?"
