"""Two-rank NCCL run of tools/check_world_invariance.py: sharded sweeps equal the unsharded ones
(bitwise for the per-session steps at fixed parameters).  Needs two GPUs on the box; on a one-GPU box
the test is skipped with that reason (the gloo tests cover the host logic there)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_two_rank_sweep_equals_one_rank(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2); profiles/r02_world_invariance_*.json holds the last run")
    out = str(tmp_path / "inv.json")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tools", "check_world_invariance.py"), out]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    res = json.loads(open(out).read())
    assert res["ok"], res
