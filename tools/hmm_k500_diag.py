"""Diagnostic: boundary mismatch of the chunked HMM filter at K = 500 (cfg5-K500 model) as a function of the
warm-up length, over a few AR-HMM sweeps; and the agreement of the chunked pass with the sequential one.
    python tools/hmm_k500_diag.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from keypointmoseq_b200 import gibbs as G, parallel
from keypointmoseq_b200.synthetic import default_hypparams, simulate

dev = torch.device("cuda")
N, T, D, L, K = 32, 31250, 10, 3, 500
pr0 = simulate(999, 1, 8, k=8, dk=2, D=D, L=L, K=K)[2]
data, st, pr, _ = simulate(1000, N, T, k=8, dk=2, D=D, L=L, K=K, params=pr0)
trans, ar, obs, loc = default_hypparams(D, L, K, full_model=False)
hyp = {}
for h in (trans, ar):
    hyp.update({a: (torch.as_tensor(b, device=dev) if isinstance(b, np.ndarray) else b) for a, b in h.items()})
d = {"mask": torch.as_tensor(data["mask"], device=dev)}
s = {"x": torch.as_tensor(st["x"], device=dev), "z": torch.as_tensor(st["z"].astype(np.int32), device=dev)}
p = {a: torch.as_tensor(pr0[a], device=dev) for a in ("Ab", "Q", "pi", "betas")}
for sweep in range(6):
    s, p = parallel.arhmm_sweep((0, sweep), d, s, p, hyp)
    pi = p["pi"]
    print("sweep", sweep, "pi: min %.3g  rows with min<1e-12: %d  diag mean %.3f  betas min %.3g  states used %d" % (
        float(pi.min()), int((pi.min(1).values < 1e-12).sum()), float(pi.diagonal().mean()), float(p["betas"].min()),
        int(torch.unique(s["z"]).numel())), flush=True)
    z0, l0, i0 = G.resample_stateseqs((9, sweep), x=s["x"], mask=d["mask"], Ab=p["Ab"], Q=p["Q"], pi=p["pi"], chunk_len=0,
                                      return_info=True)
    for burn in (256, 512, 1024, 2048, 4096):
        z1, l1, i1 = G.resample_stateseqs((9, sweep), x=s["x"], mask=d["mask"], Ab=p["Ab"], Q=p["Q"], pi=p["pi"],
                                          chunk_len=max(256, burn), burn_in=burn, strict=False, return_info=True)
        print("   burn %4d  boundary_err %.3g  z agreement %.6f  max rel dloglik %.3g" % (
            burn, float(i1["boundary_err"]), float((z0 == z1).float().mean()),
            float(((l0 - l1) / l0).abs().max())), flush=True)
print("fallbacks", G.FALLBACKS)
