set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_reference_pins.py -q -m gpu -k "ar_loglik or alternative or stateseqs or reference" > gpurun_out/r02_pytest26.log 2>&1; tail -4 gpurun_out/r02_pytest26.log
for c in cfg4-shard cfg3; do
timeout 500 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench_o_$c.json 2> gpurun_out/r02_bench_o_$c.err; echo "bench rc=$?"
done
python - <<'P'
import json
for c in ('cfg4-shard','cfg3'):
    d=json.loads(open('gpurun_out/r02_bench_o_%s.json'%c).read().strip().splitlines()[-1])
    cf=d['config']
    print(c, d['ms_per_step'], d['value'], cf['chunk_fallbacks_timed'], cf['chunk_fallbacks_all'], cf['kalman_burn_boost'], cf['hmm_burn_boost']); print(d['breakdown_ms'])
P
