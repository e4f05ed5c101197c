set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest35.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest35.log
tail -5 gpurun_out/r02_pytest35.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
for rm in 0 1; do
KPMS_HMM_LL_ROWMAJOR=$rm timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/tmp_rm$rm.json 2> gpurun_out/tmp_rm$rm.err; echo "bench rc=$?"
python - $rm <<'P'
import json,sys
d=json.loads(open('gpurun_out/tmp_rm%s.json'%sys.argv[1]).read().strip().splitlines()[-1])
print('ROWMAJOR', sys.argv[1], d['ms_per_step'], d['breakdown_ms']['stateseqs'], d['breakdown_ms']['hmm_only'], d['config']['chunk_fallbacks_all'])
P
done
KPMS_HMM_LL_ROWMAJOR=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sweeps.py -m gpu -q -k "hmm or stateseqs or sweep" > gpurun_out/r02_pytest36_rm.log 2>&1; tail -3 gpurun_out/r02_pytest36_rm.log
