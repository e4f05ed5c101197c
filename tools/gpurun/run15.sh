set -x
cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sweeps.py -q -k "latents or kalman or alternative or sweep" > gpurun_out/r02_pytest15.log 2>&1; tail -5 gpurun_out/r02_pytest15.log
timeout 900 python tools/kalman_kernels.py > gpurun_out/r02_kalman_variants3.txt 2>&1; echo "variants rc=$?"
grep -v "^$\|Warn\|_warn" gpurun_out/r02_kalman_variants3.txt | head -12
KPMS_KALMAN_OBS_MMA=0 timeout 300 python tools/kalman_kernels.py 125 36000 2>&1 | grep -v "^$\|Warn\|_warn" | head -7
