# Round 2 GPU call (1 GPU): new tests (dilute gas, special bonds through the C ABI) and the chain step after the branch-free partner test
mkdir -p gpurun_out; rm -f gpurun_out/r2j15_*
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2j15_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j15_pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 --cells 80 --no-eam --chain 8 --no-cpu --no-e2e --no-parity > gpurun_out/r2j15_chain.log 2>&1 && tail -1 gpurun_out/r2j15_chain.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); c=l['chain']; print('chain', c['value'], c['ms_per_step'], c['kernel_ms'])"
