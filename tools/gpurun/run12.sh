set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest12.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest12.log
tail -12 gpurun_out/r02_pytest12.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_g_cfg4.json 2> gpurun_out/r02_bench_g_cfg4.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_g_cfg4.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_g_cfg4.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['e2e_resident']); print(d['breakdown_ms']); print(d['sweep_timeline'])
P
CMD="python tools/sweep_once.py 125 36000 2"
$CMD > gpurun_out/r02_sweep_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_ncu_launches.csv $CMD > gpurun_out/r02_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/r02_sweep_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'kalman_forward_blk|ar_loglik_tc|hmm_forward_tc|hmm_stay_tc|kalman_backward|kalman_obs|scales_stats|heading_kernel|location_stats|hmm_ll_tile' -c 10 -f -o gpurun_out/r02_sweep_full $CMD > gpurun_out/r02_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/r02_ncu_full.log
