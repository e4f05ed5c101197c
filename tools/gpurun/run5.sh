cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x -k "latents or alternative or reference or sweep or smoke or golden" > gpurun_out/r02_pytest5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest5.log
tail -15 gpurun_out/r02_pytest5.log
timeout 600 python tools/kalman_kernels.py > gpurun_out/r02_kalman_kernels.txt 2>&1
cat gpurun_out/r02_kalman_kernels.txt | grep -v Warn | tail -24
