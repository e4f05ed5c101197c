cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x -k "latents or alternative or reference or smoke or golden" > gpurun_out/r02_pytest7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest7.log
tail -5 gpurun_out/r02_pytest7.log
timeout 600 python tools/kalman_kernels.py > gpurun_out/r02_kalman_kernels.txt 2>&1
cat gpurun_out/r02_kalman_kernels.txt | grep -v Warn | grep -E "impl|forward|max"
