set -x
cd $GRAFT_REPO_ROOT
for g in 1 3 4 6; do
timeout 500 python bench.py --e2e-groups $g --steps 5 --warmup 3 --no-cpu-baseline --no-breakdown > gpurun_out/tmp_g.json 2> gpurun_out/tmp_g.err; echo "bench rc=$?"
python - $g <<'P'
import json,sys
d=json.loads(open('gpurun_out/tmp_g.json').read().strip().splitlines()[-1])
print('GROUPS', sys.argv[1], d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'resident', d['e2e_resident']['ms_per_step'], d['config']['chunk_fallbacks_all'])
P
done 2>&1 | grep "GROUPS\|rc=[1-9]" | tee gpurun_out/r02_e2e_groups_1gpu.txt
