set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tools/check_world_invariance.py gpurun_out/r02_world_invariance_2gpu.json > gpurun_out/r02_inv2.log 2>&1; echo "inv rc=$?"; tail -3 gpurun_out/r02_inv2.log
timeout 300 python -m pytest tests/test_multi_gpu.py tests/test_gpu_sweeps.py -m gpu -q -k "two_rank or heading or ahead or reuses" > gpurun_out/r02_pytest10.log 2>&1; tail -5 gpurun_out/r02_pytest10.log
timeout 600 $TR --master-port 29534 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_e_2gpu.json 2> gpurun_out/r02_bench_e_2gpu.err; echo "bench2 rc=$?"; tail -3 gpurun_out/r02_bench_e_2gpu.err
timeout 300 $TR --master-port 29535 tools/pcie_probe_ranks.py gpurun_out/r02_pcie_2ranks.json > gpurun_out/r02_pcie2.log 2>&1; tail -2 gpurun_out/r02_pcie2.log | cut -c1-400
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_e_2gpu.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['e2e_resident']); print(d['config']['per_rank'])
P
