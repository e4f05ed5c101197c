set -x
cd $GRAFT_REPO_ROOT
for c in cfg1 cfg2 cfg3 cfg5-K20 cfg5-K100 cfg5-K500; do
  timeout 500 python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/r02_bench_h_$c.json 2> gpurun_out/r02_bench_h_$c.err; echo "$c rc=$?"
done
timeout 500 python bench.py --cold-start --steps 5 --warmup 3 > gpurun_out/r02_bench_h_cfg4_cold.json 2> gpurun_out/r02_bench_h_cfg4_cold.err; echo "cold rc=$?"
( time timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02_bench_h_reference.json 2> gpurun_out/r02_bench_h_reference.err ) 2>&1 | tail -4
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_bench_h_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('_h_')[1], d.get('ms_per_step'), d.get('value'), (d.get('e2e') or {}).get('value'), (d.get('cpu_baseline') or {}).get('value'), d['config'].get('chunk_fallbacks_timed'))
    except Exception as e:
        print(f, 'ERR', e)
P
