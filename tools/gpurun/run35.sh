set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest35.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest35.log
tail -5 gpurun_out/r02_pytest35.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/r02_final_cfg4-shard.json 2> gpurun_out/r02_final_cfg4-shard.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_final_cfg4-shard.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['cpu_baseline']['value'], d['config']['chunk_fallbacks_all'], d['roofline']['frac'], d['gpu_launches'])
P
