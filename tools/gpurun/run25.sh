set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest25.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest25.log
tail -6 gpurun_out/r02_pytest25.log
timeout 300 python tools/kalman_kernels.py 125 36000 2>&1 | grep -v "^$\|Warn\|_warn" | head -6
for c in cfg4-shard cfg3 cfg1; do
timeout 500 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_n_$c.json 2> gpurun_out/r02_bench_n_$c.err; echo "bench rc=$?"
done
python - <<'P'
import json
for c in ('cfg4-shard','cfg3','cfg1'):
    d=json.loads(open('gpurun_out/r02_bench_n_%s.json'%c).read().strip().splitlines()[-1])
    cf=d['config']
    print(c, d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e_resident']['ms_per_step'], cf['chunk_fallbacks_timed'], cf['chunk_fallbacks_all'], cf['kalman_burn_boost'], cf['hmm_burn_boost']); print(d['breakdown_ms'])
P
