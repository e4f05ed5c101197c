set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest19.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest19.log
tail -8 gpurun_out/r02_pytest19.log
timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_l_cfg4.json 2> gpurun_out/r02_bench_l_cfg4.err; echo "bench rc=$?"
python - <<'P'
import json
for f in ('gpurun_out/r02_bench_l_cfg4.json',):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e_resident']['ms_per_step']); print(d['breakdown_ms'])
P
