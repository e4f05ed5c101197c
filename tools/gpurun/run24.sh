set -x
cd $GRAFT_REPO_ROOT
for b in 64 128; do
for c in "" "--cold-start" "--config cfg3" "--config cfg5-K100" "--config cfg5-K20" "--config cfg2"; do
  timeout 400 python bench.py $c --burn-in 128 --hmm-burn-in $b --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/tmp_burn.json 2> gpurun_out/tmp_burn.err; echo "rc=$?"
  python - "$b" "$c" <<'P'
import json,sys
d=json.loads(open('gpurun_out/tmp_burn.json').read().strip().splitlines()[-1])
c=d['config']
print('HMMBURN', sys.argv[1], sys.argv[2], 'ms', round(d['ms_per_step'],2), 'fallbacks', c['chunk_fallbacks_all'], 'boost', c.get('hmm_burn_boost'), c.get('kalman_burn_boost'), 'hmm_only', round(d['breakdown_ms'].get('hmm_only',0),2), 'stateseqs', round(d['breakdown_ms'].get('stateseqs',0),2))
P
done
done 2>&1 | grep "HMMBURN\|rc=[1-9]" | tee gpurun_out/r02_hmm_burn_explore.txt
