set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sweeps.py -q -m gpu -k "host or sweep" > gpurun_out/r02_pytest38.log 2>&1; tail -6 gpurun_out/r02_pytest38.log
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-breakdown > gpurun_out/r02_bench_r_cfg4.json 2> gpurun_out/r02_bench_r_cfg4.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_r_cfg4.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_r_cfg4.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e'], d['e2e_resident'])
P
