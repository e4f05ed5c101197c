set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
./tools/microbench/fp64_probe > gpurun_out/r02_fp64_probe.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=12 -x --deselect tests/test_zz_no_alloc.py > gpurun_out/r02_pytest1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest1.log
tail -40 gpurun_out/r02_pytest1.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_a_cfg4.json 2> gpurun_out/r02_bench_a_cfg4.err; echo "bench rc=$?"
tail -3 gpurun_out/r02_bench_a_cfg4.err
for c in cfg1 cfg2 cfg5-K20 cfg5-K100 cfg5-K500 cfg3; do
  timeout 500 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/r02_bench_a_$c.json 2> gpurun_out/r02_bench_a_$c.err; echo "bench $c rc=$?"; tail -2 gpurun_out/r02_bench_a_$c.err
done
