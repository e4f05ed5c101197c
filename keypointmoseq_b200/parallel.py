"""One blocked Gibbs sweep over session shards (one process per GPU).

Sessions are conditionally independent given the parameters, so every per-frame /
per-session step runs on its own shard with no communication.  The three parameter steps
consume only sufficient statistics of the previous sweep's state, which are packed into
one float64 buffer and combined with ONE all-reduce per sweep; every rank then draws
identical parameters from the same seed (no broadcast).

Sweep order = examples/keypoint_slds.ipynb cell 12 (steps 1-8):
  ar_params, obs_variance, hdp_transitions, stateseqs, latents, heading, location, scales.
"""
import torch

from . import gibbs as G


def _world(group):
    import torch.distributed as td
    if td.is_available() and td.is_initialized():
        return td.get_rank(group), td.get_world_size(group)
    return 0, 1


def shard_range(n_sessions, rank, world):
    """Contiguous block of sessions owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_sessions, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def pack_stats(F, S_y, msum, counts):
    """-> flat float64 buffer [F | S_y | mask_sum | counts]."""
    return torch.cat([F.reshape(-1), S_y.reshape(-1), msum.reshape(-1),
                      counts.to(torch.float64).reshape(-1)])


def unpack_stats(buf, K, P, k):
    o = 0
    F = buf[o:o + K * P * P].reshape(K, P, P); o += K * P * P
    S_y = buf[o:o + k]; o += k
    msum = buf[o:o + 1]; o += 1
    counts = buf[o:o + K * K].reshape(K, K).round().to(torch.int64)
    return F, S_y, msum, counts


def sufficient_stats(data, states, hyp, Gamma, params):
    """Local (per-shard) statistics feeding the three parameter updates."""
    F = G.ar_sufficient_stats(nlags=hyp["nlags"], num_states=hyp["num_states"], mask=data["mask"],
                              x=states["x"], z=states["z"])
    S_y, msum = G.obs_variance_stats(Y=data["Y"], mask=data["mask"], Cd=params["Cd"], v=states["v"],
                                     h=states["h"], x=states["x"], s=states["s"], Gamma=Gamma)
    counts = G.transition_counts(z=states["z"], mask=data["mask"], num_states=hyp["num_states"])
    return F, S_y, msum, counts


def gibbs_sweep(key, data, states, params, hyp, Gamma=None, group=None, n_sessions_global=None):
    """One full sweep.  `data`/`states` hold this rank's shard; `params` are replicated.
    Returns new (states, params) dicts; inputs are not mutated."""
    import torch.distributed as td
    rank, world = _world(group)
    st, pr = dict(states), dict(params)
    K = int(hyp["num_states"])
    k = data["Y"].shape[2]
    N, Tm = st["z"].shape
    F, S_y, msum, counts = sufficient_stats(data, st, hyp, Gamma, pr)
    if world > 1:
        buf = pack_stats(F, S_y, msum, counts)
        td.all_reduce(buf, op=td.ReduceOp.SUM, group=group)
        F, S_y, msum, counts = unpack_stats(buf, K, F.shape[1], k)
    # 1-3: replicated parameter draws (same key on every rank)
    pr["Ab"], pr["Q"] = G.sample_ar_params(key, F, nlags=hyp["nlags"], nu_0=hyp["nu_0"], S_0=hyp["S_0"],
                                           M_0=hyp["M_0"], K_0=hyp["K_0"])
    pr["sigmasq"] = G.sample_obs_variance(key, S_y, msum, nu_sigma=hyp["nu_sigma"], sigmasq_0=hyp["sigmasq_0"])
    n_glob = n_sessions_global if n_sessions_global is not None else N * world
    pr["betas"], pr["pi"] = G.sample_hdp_transitions(key, counts, pr["betas"], hyp["alpha"], hyp["kappa"],
                                                     hyp["gamma"], max_customers=n_glob * max(Tm - 1, 1))
    # 4-8: per-session steps on the local shard; the Philox key is salted with the rank
    lkey = (key, "rank", rank)
    st["z"] = G.resample_stateseqs(lkey, x=st["x"], mask=data["mask"], Ab=pr["Ab"], Q=pr["Q"], pi=pr["pi"])[0]
    st["x"] = G.resample_latents(lkey, Y=data["Y"], mask=data["mask"], v=st["v"], h=st["h"], z=st["z"],
                                 s=st["s"], Cd=pr["Cd"], sigmasq=pr["sigmasq"], Ab=pr["Ab"], Q=pr["Q"],
                                 Gamma=Gamma)
    st["h"] = G.resample_heading(lkey, Y=data["Y"], v=st["v"], x=st["x"], s=st["s"], Cd=pr["Cd"],
                                 sigmasq=pr["sigmasq"], Gamma=Gamma)
    st["v"] = G.resample_location(lkey, mask=data["mask"], Y=data["Y"], h=st["h"], x=st["x"], s=st["s"],
                                  Cd=pr["Cd"], sigmasq=pr["sigmasq"], sigmasq_loc=hyp["sigmasq_loc"],
                                  Gamma=Gamma)
    st["s"] = G.resample_scales(lkey, x=st["x"], v=st["v"], h=st["h"], Y=data["Y"], Cd=pr["Cd"],
                                sigmasq=pr["sigmasq"], nu_s=hyp["nu_s"], s_0=hyp["s_0"], Gamma=Gamma)
    return st, pr
