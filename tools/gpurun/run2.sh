set -x
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -q --maxfail=8 > gpurun_out/r02_pytest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest2.log
tail -60 gpurun_out/r02_pytest2.log
for c in cfg5-K500 cfg3 cfg1; do
  timeout 500 python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench_b_$c.json 2> gpurun_out/r02_bench_b_$c.err; echo "bench $c rc=$?"; tail -2 gpurun_out/r02_bench_b_$c.err
done
