"""Explore chunk_len / burn_in of the chunk-parallel Kalman path on the bench workload:
prints time per call, boundary mismatch, and max |x - x_exact|."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from keypointmoseq_b200 import gibbs as G
from keypointmoseq_b200.synthetic import simulate

N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 36000
ragged = len(sys.argv) > 3 and sys.argv[3] == "ragged"
k, dk, D, L, K = 16, 2, 10, 3, 100
data, st, pr, Gamma = simulate(5, N, T, k=k, dk=dk, D=D, L=L, K=K, ragged=ragged)
dev = "cuda"
tt = lambda a: torch.as_tensor(a, device=dev)
d = {a: tt(b) for a, b in data.items()}
s = {a: tt(b.astype(np.int32) if a == 'z' else b) for a, b in st.items()}
p = {a: tt(b) for a, b in pr.items()}
w = torch.randn((N, T - L, D * L), device=dev); wl = torch.randn((N, D * L), device=dev)
kw = dict(**d, **s, **p, Gamma=tt(Gamma), draws={'w': w, 'w_last': wl}, return_info=True)

def run(cl, bi, reps=2):
    x, info = G.resample_latents(0, chunk_len=cl, burn_in=bi, **kw)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        x, info = G.resample_latents(0, chunk_len=cl, burn_in=bi, **kw)
    b.record(); torch.cuda.synchronize()
    return x, info, a.elapsed_time(b) / reps

x0, i0, t0 = run(0, 0, reps=1)
print(json.dumps(dict(mode="exact", ms=t0, frames=N * T)), flush=True)
pairs = [(2048, 256), (1024, 256), (512, 256), (512, 128), (256, 128), (256, 64), (128, 64)]
if len(sys.argv) > 4:
    pairs = [tuple(int(v) for v in a.split(':')) for a in sys.argv[4].split(',')]
for cl, bi in pairs:
    x, info, t = run(cl, bi)
    be = info['boundary_err'].cpu().numpy()
    dx = float((x - x0).abs().max())
    dl = float(((info['loglik'] - i0['loglik']).abs() / i0['loglik'].abs()).max())
    print(json.dumps(dict(chunk_len=cl, burn_in=bi, ms=t, berr_filter=float(be[0]), berr_sampler=float(be[1]),
                          max_dx=dx, max_rel_dll=dl)), flush=True)
