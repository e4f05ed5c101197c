"""One device-resident sweep on the bench shard (for ncu captures)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from keypointmoseq_b200 import gibbs as G, parallel
from keypointmoseq_b200.synthetic import simulate, default_hypparams
N = int(sys.argv[1]) if len(sys.argv) > 1 else 125
T = int(sys.argv[2]) if len(sys.argv) > 2 else 36000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
k, dk, D, L, K = 16, 2, 10, 3, 100
pr = simulate(999, 1, 8, k=k, dk=dk, D=D, L=L, K=K)[2]          # as bench.py: the model the shard is drawn from
data, st, _, Gamma = simulate(1000, N, T, k=k, dk=dk, D=D, L=L, K=K, params=pr)
dev = torch.device("cuda")
tt = lambda a: torch.as_tensor(a, device=dev)
d = {a: tt(b) for a, b in data.items()}
s = {a: tt(b.astype(np.int32) if a == 'z' else b) for a, b in st.items()}
p = {a: tt(b) for a, b in pr.items()}
hyp = {}
for h in default_hypparams(D, L, K):
    hyp.update({a: (tt(b) if isinstance(b, np.ndarray) else b) for a, b in h.items()})
for i in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    s, p = parallel.gibbs_sweep((0, i), d, s, p, hyp, Gamma=tt(Gamma), next_key=(0, i + 1))
    b.record()
    torch.cuda.synchronize()
    print(i, "ms", round(a.elapsed_time(b), 2), "hmm berr", float(G.last_stateseq_boundary_err), "kalman berr",
          G.last_latents_boundary_err.tolist(), dict(G.FALLBACKS), flush=True)
torch.cuda.synchronize()
print("ok", float(s['x'].abs().mean()))
