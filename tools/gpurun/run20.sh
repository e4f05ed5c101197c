set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_kalman_scan.py tests/test_gpu_sweeps.py -q -m gpu > gpurun_out/r02_pytest20.log 2>&1; tail -5 gpurun_out/r02_pytest20.log
timeout 500 python bench.py --config cfg5-K500 --steps 12 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_m_cfg5-K500.json 2> gpurun_out/r02_bench_m_cfg5-K500.err; echo "rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_m_cfg5-K500.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['config']['chunk_fallbacks_timed'], d['config']['chunk_fallbacks_all'], d['config']['hmm_burn_boost'])
P
