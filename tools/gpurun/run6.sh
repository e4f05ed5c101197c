cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x -k "latents or alternative or reference or sweep or smoke or golden" > gpurun_out/r02_pytest6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest6.log
tail -8 gpurun_out/r02_pytest6.log
timeout 600 python tools/kalman_kernels.py > gpurun_out/r02_kalman_kernels.txt 2>&1
cat gpurun_out/r02_kalman_kernels.txt | grep -v Warn | tail -22
timeout 900 ncu --set full --clock-control none --import-source on -k regex:kalman_forward_blk --launch-count 1 -o gpurun_out/r02_kalman_blk_v2 -f python tools/kalman_once.py 125 36000 1 > gpurun_out/r02_ncu_kalman.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/r02_ncu_kalman.log
