set -x
cd $GRAFT_REPO_ROOT
export KPMS_SAN=1
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "resample_latents_parity or ar_loglik_parity or repaired_large_K or host_sweep_session_groups or scales or heading" > gpurun_out/r02_memcheck.log 2>&1; echo "memcheck rc=$?"
grep -c "Invalid\|out of bounds\|misaligned" gpurun_out/r02_memcheck.log; tail -6 gpurun_out/r02_memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 --print-limit 20 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "resample_latents_parity and 16-10" > gpurun_out/r02_racecheck.log 2>&1; echo "racecheck rc=$?"
grep -c "hazard" gpurun_out/r02_racecheck.log; tail -6 gpurun_out/r02_racecheck.log
