set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x -k "latents or alternative or reference or sweep or smoke or golden or notebook or fit" > gpurun_out/r02_pytest3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest3.log
tail -40 gpurun_out/r02_pytest3.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench_c_cfg4.json 2> gpurun_out/r02_bench_c_cfg4.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_c_cfg4.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r02_bench_c_cfg4.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['breakdown_ms'], d['config']['chunk_fallbacks_timed'], d['config']['kalman_boundary_err'])
P
