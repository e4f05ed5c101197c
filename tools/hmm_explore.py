"""Time the HMM forward-filter/backward-sampler kernel alone (K=100 by default)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from keypointmoseq_b200 import gibbs as G
N = int(sys.argv[1]) if len(sys.argv) > 1 else 125
T = int(sys.argv[2]) if len(sys.argv) > 2 else 36000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 100
g = torch.Generator(device="cuda"); g.manual_seed(0)
ll = torch.randn((N, T, K), device="cuda", generator=g) * 3 - 4
mask = torch.ones((N, T), device="cuda")
pi = torch.rand((K, K), device="cuda", generator=g, dtype=torch.float64) * 0.1 + 0.9 * torch.eye(K, device="cuda", dtype=torch.float64)
pi = pi / pi.sum(1, keepdim=True)
u = torch.rand((N, T), device="cuda", generator=g)
for rep in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); z, lg = G.sample_hmm_stateseqs(ll=ll, mask=mask, pi=pi, u=u); b.record(); torch.cuda.synchronize()
    print(json.dumps(dict(N=N, T=T, K=K, ms=a.elapsed_time(b), us_per_step=a.elapsed_time(b) * 1e3 / T)), flush=True)
