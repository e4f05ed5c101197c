"""One chunked resample_latents call on the bench shard (for ncu captures of the Kalman kernels)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from keypointmoseq_b200 import gibbs as G
from keypointmoseq_b200.synthetic import simulate
N = int(sys.argv[1]) if len(sys.argv) > 1 else 125
T = int(sys.argv[2]) if len(sys.argv) > 2 else 36000
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
k, dk, D, L, K = 16, 2, 10, 3, 100
data, st, pr, Gamma = simulate(5, N, T, k=k, dk=dk, D=D, L=L, K=K)
tt = lambda a: torch.as_tensor(a, device="cuda")
d = {a: tt(b) for a, b in data.items()}
s = {a: tt(b.astype(np.int32) if a == 'z' else b) for a, b in st.items()}
p = {a: tt(b) for a, b in pr.items()}
w = torch.randn((N, T - L, D * L), device="cuda"); wl = torch.randn((N, D * L), device="cuda")
kw = dict(**d, **s, **p, Gamma=tt(Gamma), draws={'w': w, 'w_last': wl}, return_info=True)
for _ in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); x, info = G.resample_latents(0, **kw); b.record(); torch.cuda.synchronize()
    print("ms", a.elapsed_time(b), info['boundary_err'].cpu().tolist(), flush=True)
