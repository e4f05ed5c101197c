set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sweeps.py -q -k "ar_loglik or alternative or stateseqs or sweep" > gpurun_out/r02_pytest17.log 2>&1; tail -5 gpurun_out/r02_pytest17.log
timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench_j_cfg4.json 2> gpurun_out/r02_bench_j_cfg4.err; echo "bench rc=$?"
KPMS_EMISSION_PIPE=0 timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench_j_cfg4_nopipe.json 2> gpurun_out/r02_bench_j_cfg4_nopipe.err; echo "bench rc=$?"
timeout 400 python bench.py --config cfg3 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench_j_cfg3.json 2> gpurun_out/r02_bench_j_cfg3.err; echo "bench rc=$?"
python - <<'P'
import json
for f in ('gpurun_out/r02_bench_j_cfg4.json','gpurun_out/r02_bench_j_cfg4_nopipe.json','gpurun_out/r02_bench_j_cfg3.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d['value']); print(d['breakdown_ms'])
P
