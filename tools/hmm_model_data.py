"""Time resample_stateseqs on model-simulated (sticky) data and report the state-change rate."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from keypointmoseq_b200 import gibbs as G
from keypointmoseq_b200.synthetic import simulate
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16
T = int(sys.argv[2]) if len(sys.argv) > 2 else 36000
data, st, pr, Gamma = simulate(5, N, T, k=8, dk=2, D=10, L=3, K=100)
tt = lambda a: torch.as_tensor(a, device="cuda")
x, mask = tt(st['x']), tt(data['mask'])
Ab, Q, pi = tt(pr['Ab']), tt(pr['Q']), tt(pr['pi'])
ll = G.ar_log_likelihood(x=x, Ab=Ab, Q=Q)
u = torch.rand((N, T - 3), device="cuda")
z0, lg0 = G.sample_hmm_stateseqs(ll=ll, mask=mask[:, 3:], pi=pi, u=u, chunk_len=0)
modes = [(0, 0), ("auto", 256), (258, 96), (258, 64), (258, 48), (258, 32), (258, 16)]
for cl, bi in modes:
    z, lg, info = G.sample_hmm_stateseqs(ll=ll, mask=mask[:, 3:], pi=pi, u=u, chunk_len=cl, burn_in=bi, strict=False, return_info=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); z, lg, info = G.sample_hmm_stateseqs(ll=ll, mask=mask[:, 3:], pi=pi, u=u, chunk_len=cl, burn_in=bi, strict=False, return_info=True); b.record(); torch.cuda.synchronize()
    print(json.dumps(dict(chunk_len=cl, burn_in=bi, berr=float(info['boundary_err']), dz=float((z != z0).float().mean()), dll=float(((lg - lg0).abs() / lg0.abs()).max()))), end=' ')
    chg = float((z[:, 1:] != z[:, :-1]).float().mean())
    agree = float((z.cpu().numpy() == st['z']).mean())
    print(json.dumps(dict(N=N, T=T, ms=a.elapsed_time(b), us_per_step=a.elapsed_time(b) * 1e3 / T, change_rate=chg, agree_with_truth=agree)), flush=True)
