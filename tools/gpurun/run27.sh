set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29546 bench.py --gpus 4 --steps 5 --warmup 3 --no-breakdown > gpurun_out/r02_final_cfg4_4gpu.json 2> gpurun_out/r02_final_cfg4_4gpu.err; echo "bench8 rc=$?"; tail -3 gpurun_out/r02_final_cfg4_4gpu.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_final_cfg4_4gpu.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['e2e_resident']); print(d['config']['per_rank']); print(d.get("breakdown_ms"))
P
