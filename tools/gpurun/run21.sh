set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest21.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest21.log
tail -6 gpurun_out/r02_pytest21.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke21.log 2>&1; tail -2 gpurun_out/r02_smoke21.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_final_cfg4-shard.json 2> gpurun_out/r02_final_cfg4-shard.err; echo "bench rc=$?"
for c in cfg1 cfg2 cfg3 cfg5-K20 cfg5-K100 cfg5-K500; do
  timeout 500 python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/r02_final_$c.json 2> gpurun_out/r02_final_$c.err; echo "$c rc=$?"
done
timeout 500 python bench.py --cold-start --steps 5 --warmup 3 > gpurun_out/r02_final_cfg4-shard_cold.json 2> gpurun_out/r02_final_cfg4-shard_cold.err; echo "cold rc=$?"
timeout 900 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02_final_reference_arm.json 2> gpurun_out/r02_final_reference_arm.err; echo "ref rc=$?"
CMD="python tools/sweep_once.py 125 36000 2"
$CMD > gpurun_out/r02_sweep_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches.csv $CMD > gpurun_out/r02_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/r02_sweep_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'kalman_forward_blk|ar_loglik_tc|hmm_forward_tc|hmm_stay_tc|kalman_backward|kalman_obs|scales_stats|heading_kernel|location_stats|hmm_ll_tile' -c 10 -f -o gpurun_out/r02_sweep_full $CMD > gpurun_out/r02_ncu_full.log 2>&1; echo "ncu full rc=$?"
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02_final_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('final_')[1], d.get('ms_per_step'), d.get('value'), (d.get('e2e') or {}).get('value'), (d.get('cpu_baseline') or {}).get('value'), d['config'].get('chunk_fallbacks_timed'))
    except Exception as e:
        print(f, 'ERR', e)
P
