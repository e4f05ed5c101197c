"""One chunked FFBS call on the bench shard (for ncu captures of the HMM kernels)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from keypointmoseq_b200 import gibbs as G
N, T, K = 125, 36000, 100
g = torch.Generator(device="cuda").manual_seed(1)
ll = torch.randn((N, T - 3, K), device="cuda", generator=g) * 3.0
pi = torch.full((K, K), 0.05 / (K - 1), device="cuda", dtype=torch.float64)
pi.fill_diagonal_(0.95)
mask = torch.ones((N, T - 3), device="cuda")
u = torch.rand((N, T - 3), device="cuda", generator=g)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for _ in range(reps):
    z, lg, info = G.sample_hmm_stateseqs(ll=ll, mask=mask, pi=pi, u=u, chunk_len="auto", burn_in=256, strict=False, return_info=True)
torch.cuda.synchronize()
print(float(info["boundary_err"]), float(lg.sum()))
