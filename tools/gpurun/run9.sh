set -x
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest9.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest9.log
tail -40 gpurun_out/r02_pytest9.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_d_cfg4.json 2> gpurun_out/r02_bench_d_cfg4.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_d_cfg4.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_d_cfg4.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['e2e_resident'])
print(d['breakdown_ms']); print(d['sweep_timeline']); print(d['config']['chunk_fallbacks_timed'])
P
