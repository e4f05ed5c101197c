"""Per-kernel device times (torch.profiler / CUPTI) of one resample_latents call on the bench shard,
for the blocked DMMA filter (kalman_impl = 2) and the lane-per-column warp filter (kalman_impl = 1).
    python tools/kalman_kernels.py [sessions] [frames]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

from keypointmoseq_b200 import _lib, gibbs as G
from keypointmoseq_b200.synthetic import simulate

N = int(sys.argv[1]) if len(sys.argv) > 1 else 125
T = int(sys.argv[2]) if len(sys.argv) > 2 else 36000
k, dk, D, L, K = 16, 2, 10, 3, 100
data, st, pr, Gamma = simulate(5, N, T, k=k, dk=dk, D=D, L=L, K=K)
tt = lambda a: torch.as_tensor(a, device="cuda")
d = {a: tt(b) for a, b in data.items()}
s = {a: tt(b.astype(np.int32) if a == 'z' else b) for a, b in st.items()}
p = {a: tt(b) for a, b in pr.items()}
w = torch.randn((N, T - L, D * L), device="cuda")
wl = torch.randn((N, D * L), device="cuda")
kw = dict(**d, **s, **p, Gamma=tt(Gamma), draws={'w': w, 'w_last': wl}, return_info=True)
out = {}
variants = [(2, 12, 0), (2, 12, 2), (2, 8, 0), (2, 8, 2), (2, 16, 0), (2, 16, 2), (2, 12, 1), (1, 12, 1)]
for impl, warps, lock in variants:
    _lib.set_option("kalman_impl", impl)
    _lib.set_option("kalman_warps", warps)
    _lib.set_option("kalman_lockstep", lock)
    G.KALMAN_WARPS_PER_SM = warps
    print("---- impl", impl, "warps/CTA", warps, "lockstep/prefetch mask", lock)
    G.resample_latents(0, **kw)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        x, info = G.resample_latents(0, **kw)
        torch.cuda.synchronize()
    out[impl] = x
    top = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:4 if (impl, warps, lock) == variants[0] else 1]
    print("== kalman_impl", impl, "boundary_err", info['boundary_err'].cpu().tolist(), "loglik[0]", float(info['loglik'][0]))
    for e in top:
        print("   %10.3f ms  x%d  %s" % (e.device_time_total / 1e3, e.count, e.key[:100]))
print("max |x(impl 2) - x(impl 1)| =", float((out[2] - out[1]).abs().max()))
