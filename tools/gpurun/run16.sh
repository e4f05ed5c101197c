set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "latents or kalman or alternative" > gpurun_out/r02_pytest16.log 2>&1; tail -3 gpurun_out/r02_pytest16.log
timeout 300 python tools/kalman_kernels.py 125 36000 2>&1 | grep -v "^$\|Warn\|_warn" | head -7
timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_i_cfg4.json 2> gpurun_out/r02_bench_i_cfg4.err; echo "bench rc=$?"
KPMS_LL_TILED=1 timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_i_cfg4_lltiled.json 2> gpurun_out/r02_bench_i_cfg4_lltiled.err; echo "bench rc=$?"
python - <<'P'
import json
for f in ('gpurun_out/r02_bench_i_cfg4.json','gpurun_out/r02_bench_i_cfg4_lltiled.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['e2e_resident']['ms_per_step']); print(d['breakdown_ms'])
P
