cd $GRAFT_REPO_ROOT
timeout 600 python tools/kalman_kernels.py > gpurun_out/r02_kalman_kernels.txt 2>&1
cat gpurun_out/r02_kalman_kernels.txt | tail -30
