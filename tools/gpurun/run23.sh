set -x
cd $GRAFT_REPO_ROOT
for b in 128 192; do
for c in "" "--cold-start" "--config cfg3" "--config cfg1"; do
  timeout 400 python bench.py $c --burn-in $b --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-breakdown > gpurun_out/tmp_burn.json 2> gpurun_out/tmp_burn.err; echo "rc=$?"
  python - "$b" "$c" <<'P'
import json,sys
d=json.loads(open('gpurun_out/tmp_burn.json').read().strip().splitlines()[-1])
c=d['config']
print('BURN', sys.argv[1], sys.argv[2], 'ms', round(d['ms_per_step'],2), 'fallbacks', c['chunk_fallbacks_all'], 'berr', c.get('kalman_boundary_err'), 'cl', c.get('kalman_chunk_len'), 'boost', c.get('kalman_burn_boost'))
P
done
done 2>&1 | grep "BURN\|rc=" | tee gpurun_out/r02_burn_explore.txt
