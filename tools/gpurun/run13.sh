set -x
cd $GRAFT_REPO_ROOT
timeout 900 python tools/kalman_kernels.py > gpurun_out/r02_kalman_variants2.txt 2>&1; echo "variants rc=$?"
grep -v "^$" gpurun_out/r02_kalman_variants2.txt | tail -30
timeout 300 python -m pytest tests/test_gpu_parity.py -q -k "latents or kalman" > gpurun_out/r02_pytest13.log 2>&1; tail -3 gpurun_out/r02_pytest13.log
