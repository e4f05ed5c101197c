set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29534 bench.py --gpus 2 --steps 5 --warmup 3 --no-breakdown > gpurun_out/r02_final_cfg4_2gpu.json 2> gpurun_out/r02_final_cfg4_2gpu.err; echo "bench2 rc=$?"; tail -2 gpurun_out/r02_final_cfg4_2gpu.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_final_cfg4_2gpu.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e_resident']['value'])
P
