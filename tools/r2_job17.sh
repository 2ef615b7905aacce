#!/bin/bash
# new no-alloc entry point alone in a fresh process, then the whole GPU suite (it runs last), then the default bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/j17_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/j17_noalloc_alone.log
tail -3 gpurun_out/j17_noalloc_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/j17_pytest.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/j17_pytest.log
tail -4 gpurun_out/j17_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/j17_bench.json 2> gpurun_out/j17_bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/j17_bench.json
