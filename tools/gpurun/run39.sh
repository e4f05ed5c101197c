set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_sweeps.py tests/test_gpu_parity.py -q -m gpu -k "sweep or host or latents" > gpurun_out/r02_pytest39.log 2>&1; tail -4 gpurun_out/r02_pytest39.log
timeout 500 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-breakdown > gpurun_out/r02_bench_s_cfg4.json 2> gpurun_out/r02_bench_s_cfg4.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_s_cfg4.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e_resident']['ms_per_step'], d['sweep_timeline'])
P
