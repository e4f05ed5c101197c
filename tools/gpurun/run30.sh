set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for g in 2 4; do
timeout 600 $TR --master-port $((29550+g)) bench.py --gpus 8 --steps 5 --warmup 3 --no-breakdown --e2e-groups $g > gpurun_out/r02_bench_q_8gpu_g$g.json 2> gpurun_out/r02_bench_q_8gpu_g$g.err; echo "bench8 rc=$?"; tail -2 gpurun_out/r02_bench_q_8gpu_g$g.err
python - $g <<'P'
import json,sys
d=json.loads(open('gpurun_out/r02_bench_q_8gpu_g%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print('GROUPS8', sys.argv[1], d['ms_per_step'], d['value'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['value'], 'resident', d['e2e_resident']['ms_per_step'], d['e2e_resident']['value'])
P
done
