# Round 2 GPU call: build kernel re-measure (1 GPU part of the 2-GPU box), then the 2-GPU decomposition tests + bench
mkdir -p gpurun_out; rm -f gpurun_out/r2j9_*
nvidia-smi -L | head -3
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2j9_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j9_pytest.log
timeout 300 python tools/kbench.py lj 160 seg > gpurun_out/r2j9_kbench_lj.log 2>&1; tail -4 gpurun_out/r2j9_kbench_lj.log | head -2
timeout 300 python tools/kbench.py eam 100 seg > gpurun_out/r2j9_kbench_eam.log 2>&1; tail -4 gpurun_out/r2j9_kbench_eam.log | head -2
timeout 300 python bench.py --steps 20 --warmup 5 --cells 80 --no-eam --chain 8 --no-cpu --no-e2e --no-parity > gpurun_out/r2j9_chain.log 2>&1 && tail -1 gpurun_out/r2j9_chain.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); c=l['chain']; print('chain', c['value'], c['ms_per_step'], c['kernel_ms'])"
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2j9_bench2.log 2>&1
echo "bench2 rc=$?"; tail -1 gpurun_out/r2j9_bench2.log | cut -c1-300
