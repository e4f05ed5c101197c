set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest33.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest33.log
tail -6 gpurun_out/r02_pytest33.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tools/check_world_invariance.py gpurun_out/r02_world_invariance_2gpu.json > gpurun_out/r02_inv2.log 2>&1; echo "inv rc=$?"; tail -1 gpurun_out/r02_inv2.log | cut -c1-500
timeout 600 $TR --master-port 29534 bench.py --gpus 2 --steps 5 --warmup 3 --no-breakdown > gpurun_out/r02_final_cfg4_2gpu.json 2> gpurun_out/r02_final_cfg4_2gpu.err; echo "bench2 rc=$?"; tail -2 gpurun_out/r02_final_cfg4_2gpu.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_final_cfg4_2gpu.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e_resident']['value']); print(d['config']['per_rank'])
P
