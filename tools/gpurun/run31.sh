set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest31.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest31.log
tail -6 gpurun_out/r02_pytest31.log
timeout 600 python bench.py --config cfg5-K500 --steps 10 --warmup 3 > gpurun_out/r02_final_cfg5-K500.json 2> gpurun_out/r02_final_cfg5-K500.err; echo "rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_final_cfg5-K500.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['chunk_fallbacks_timed'], d['config']['chunk_fallbacks_all'], d['config']['hmm_burn_boost'])
P
