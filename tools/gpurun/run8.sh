cd $GRAFT_REPO_ROOT
timeout 900 python tools/kalman_kernels.py > gpurun_out/r02_kalman_variants.txt 2>&1
grep -v Warn gpurun_out/r02_kalman_variants.txt | grep -E "^----|forward|max"
