"""Where the time of one host-driven sweep goes (ms since entry, main-stream events)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from keypointmoseq_b200 import parallel
from keypointmoseq_b200.synthetic import simulate, default_hypparams
N, T, k, dk, D, L, K = 125, 36000, 16, 2, 10, 3, 100
data, st, pr, Gamma = simulate(1000, N, T, k=k, dk=dk, D=D, L=L, K=K)
dev = torch.device("cuda")
hyp = {}
for h in default_hypparams(D, L, K):
    hyp.update({a: (torch.as_tensor(b, device=dev) if isinstance(b, np.ndarray) else b) for a, b in h.items()})
host = {a: torch.as_tensor(b).pin_memory() for a, b in data.items()}
host_st = {a: torch.as_tensor(b.astype(np.int32) if a == "z" else b).pin_memory() for a, b in st.items()}
p = {a: torch.as_tensor(b, device=dev) for a, b in pr.items()}
Gm = torch.as_tensor(Gamma, device=dev)
out = None
for i in range(5):
    tl = {}
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out, p, _ = parallel.gibbs_sweep_host((0, i), host, host_st, p, hyp, Gamma=Gm, out=out, timeline=tl)
    torch.cuda.synchronize(); wall = (time.perf_counter() - t0) * 1e3
    from keypointmoseq_b200 import gibbs as G
    print("wall %.1f ms " % wall + " ".join("%s=%.1f" % kv for kv in tl.items()), dict(G.FALLBACKS), G._HMM_ADAPT, flush=True)
