"""Sharded = unsharded: run under torchrun with G ranks (one per GPU, NCCL).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tools/check_world_invariance.py [out.json]

A fixed problem of N sessions (N not divisible by G) is sharded contiguously over the ranks.
 (a) per-session steps with FIXED parameters and session_offset = first session of the shard must
     reproduce, bit for bit, the rows of the same steps run unsharded (rank 0 runs those on all N
     sessions): the Philox draws are keyed by the global session index.
 (b) two full `parallel.gibbs_sweep` calls (packed all-reduce of the statistics, replicated parameter
     draws): parameters identical on all ranks and equal to the unsharded sweep's up to the rounding of
     the differently ordered sums; state sequences >= 99.9 % equal, latents close.
Rank 0 prints one JSON line (and writes it to out.json) with "ok": true/false.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as td

from keypointmoseq_b200 import gibbs as G, parallel
from keypointmoseq_b200.synthetic import default_hypparams, simulate


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    td.init_process_group("nccl", device_id=dev)
    solo = td.new_group([0])          # a one-rank group: the unsharded run, on rank 0
    N, T, k, D, L, K = 7, 3000, 8, 10, 3, 12
    data, st, pr, Gamma = simulate(314, N, T, k=k, dk=2, D=D, L=L, K=K, ragged=True)
    trans, ar, obs, loc = default_hypparams(D, L, K)
    hyp = {}
    for h in (trans, ar, obs, loc):
        hyp.update({a: (torch.as_tensor(b, device=dev) if isinstance(b, np.ndarray) else b) for a, b in h.items()})
    tt = lambda a: torch.as_tensor(a, device=dev)
    full_d = {a: tt(b) for a, b in data.items()}
    full_s = {a: tt(b.astype(np.int32) if a == "z" else b) for a, b in st.items()}
    p = {a: tt(b) for a, b in pr.items()}
    Gm = tt(Gamma)
    lo, hi = parallel.shard_range(N, rank, world)
    d = {a: b[lo:hi].contiguous() for a, b in full_d.items()}
    s = {a: b[lo:hi].contiguous() for a, b in full_s.items()}
    key = (7, 1)
    res = {"world": world, "sessions": N, "shard": [lo, hi]}

    def steps(dd, ss, off):
        kw = dict(Gamma=Gm, session_offset=off)
        o = {}
        o["z"] = G.resample_stateseqs(key, **dd, **ss, **p, chunk_len=500, burn_in=256, session_offset=off)[0]
        o["x"] = G.resample_latents(key, **dd, **ss, **p, chunk_len=600, burn_in=256, **kw)
        o["h"] = G.resample_heading(key, **dd, **ss, **p, **kw)
        o["v"] = G.resample_location(key, **dd, **ss, **p, sigmasq_loc=hyp["sigmasq_loc"], **kw)
        o["s"] = G.resample_scales(key, **dd, **ss, **p, nu_s=hyp["nu_s"], s_0=hyp["s_0"], **kw)
        return o

    # (a) fixed parameters: bitwise
    mine = steps(d, s, lo)
    whole = steps(full_d, full_s, 0)      # every rank can afford it at this size; compared locally
    res_a = {name: bool(torch.equal(mine[name], whole[name][lo:hi])) for name in mine}
    flags = torch.tensor([float(all(res_a.values()))], device=dev)
    td.all_reduce(flags, op=td.ReduceOp.MIN)
    res["fixed_params_bitwise_equal"] = bool(flags.item() > 0)
    res["fixed_params_detail_rank0"] = res_a

    # (b) two full sweeps, sharded vs unsharded
    ss, pp = s, p
    for i in range(2):
        ss, pp = parallel.gibbs_sweep((3, i), d, ss, pp, hyp, Gamma=Gm, session_offset=lo)
    par = torch.cat([pp[a].reshape(-1) for a in ("Ab", "Q", "pi", "betas", "sigmasq")])
    gathered = [torch.empty_like(par) for _ in range(world)]
    td.all_gather(gathered, par)
    res["params_identical_across_ranks"] = bool(all(torch.equal(g, par) for g in gathered))
    zs, xs = [None] * world, [None] * world     # shards differ in size: gather as objects
    td.all_gather_object(zs, ss["z"].cpu())
    td.all_gather_object(xs, ss["x"].cpu())
    if rank == 0:
        us, up = full_s, p
        for i in range(2):
            us, up = parallel.gibbs_sweep((3, i), full_d, us, up, hyp, Gamma=Gm, group=solo, session_offset=0)
        res["params_max_rel_diff_vs_unsharded"] = max(
            float((pp[a] - up[a]).abs().max() / up[a].abs().max()) for a in ("Ab", "Q", "pi", "betas", "sigmasq"))
        zcat, xcat = torch.cat(zs).to(dev), torch.cat(xs).to(dev)
        valid = full_d["mask"][:, L:] > 0
        res["z_agreement_vs_unsharded"] = float((zcat[valid] == us["z"][valid]).float().mean())
        res["x_max_abs_diff_vs_unsharded"] = float((xcat - us["x"]).abs().max())
        res["ok"] = bool(res["fixed_params_bitwise_equal"] and res["params_identical_across_ranks"] and
                         res["params_max_rel_diff_vs_unsharded"] < 1e-5 and res["z_agreement_vs_unsharded"] >= 0.999)
        line = json.dumps(res)
        print(line, flush=True)
        if len(sys.argv) > 1:
            with open(sys.argv[1], "w") as fh:
                fh.write(line + "\n")
    td.barrier()
    td.destroy_process_group()


if __name__ == "__main__":
    main()
