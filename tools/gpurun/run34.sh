set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sweeps.py -q -m gpu -k "hmm or stateseqs or K500 or large_K or k500 or 500" > gpurun_out/r02_pytest34.log 2>&1; tail -8 gpurun_out/r02_pytest34.log
timeout 600 python bench.py --config cfg5-K500 --steps 10 --warmup 3 > gpurun_out/r02_final_cfg5-K500.json 2> gpurun_out/r02_final_cfg5-K500.err; echo "rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_final_cfg5-K500.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['config']['chunk_fallbacks_timed'], d['config']['chunk_fallbacks_all'], d['config']['hmm_burn_boost']); print(d['breakdown_ms'])
P
