cd $GRAFT_REPO_ROOT
timeout 900 python tools/hmm_k500_diag.py > gpurun_out/r02_hmm_k500_diag.txt 2>&1; echo rc=$?; cat gpurun_out/r02_hmm_k500_diag.txt | tail -45
