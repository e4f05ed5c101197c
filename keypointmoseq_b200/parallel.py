"""One blocked Gibbs sweep over session shards (one process per GPU).

Sessions are conditionally independent given the parameters, so every per-frame /
per-session step runs on its own shard with no communication.  The three parameter steps
consume only sufficient statistics of the previous sweep's state, which are packed into
one float64 buffer and combined with ONE all-reduce per sweep; every rank then draws
identical parameters from the same seed (no broadcast).

Sweep order = examples/keypoint_slds.ipynb cell 12 (steps 1-8):
  ar_params, obs_variance, hdp_transitions, stateseqs, latents, heading, location, scales.
"""
import os

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


def all_reduce_stats(F, S_y, msum, counts, group=None):
    """Sum the per-shard statistics over the ranks with ONE all-reduce of the packed buffer.
    S_y / msum may be None (they are then reduced later by `all_reduce_obs_stats`)."""
    import torch.distributed as td
    K, P, k = counts.shape[0], F.shape[1], (S_y.numel() if S_y is not None else 1)
    buf = pack_stats(F, S_y if S_y is not None else F.new_zeros(k), msum if msum is not None else F.new_zeros(1),
                     counts)
    td.all_reduce(buf, op=td.ReduceOp.SUM, group=group)
    F, S2, m2, counts = unpack_stats(buf, K, P, k)
    return F, (S2 if S_y is not None else None), (m2 if msum is not None else None), counts


def all_reduce_obs_stats(S_y, msum, group=None):
    """The deferred observation-noise statistics of `gibbs_sweep_host`: k + 1 doubles."""
    import torch.distributed as td
    k = S_y.numel()
    buf = torch.cat([S_y.reshape(-1), msum.reshape(-1)])
    td.all_reduce(buf, op=td.ReduceOp.SUM, group=group)
    return buf[:k], buf[k:k + 1]


def sufficient_stats(data, states, hyp, Gamma, params):
    """Local (per-shard) statistics feeding the three parameter updates."""
    F = G.ar_sufficient_stats(nlags=hyp["nlags"], num_states=hyp["num_states"], mask=data["mask"],
                              x=states["x"], z=states["z"])
    S_y, msum = G.obs_variance_stats(Y=data["Y"], mask=data["mask"], Cd=params["Cd"], v=states["v"],
                                     h=states["h"], x=states["x"], s=states["s"], Gamma=Gamma)
    counts = G.transition_counts(z=states["z"], mask=data["mask"], num_states=hyp["num_states"])
    return F, S_y, msum, counts


def gibbs_sweep(key, data, states, params, hyp, Gamma=None, group=None, n_sessions_global=None,
                _wait=None, _done=None, _mark=None):
    """One full sweep.  `data`/`states` hold this rank's shard; `params` are replicated.
    Returns new (states, params) dicts; inputs are not mutated.

    `_wait(group)` / `_done(name, tensor)` are the hooks of `gibbs_sweep_host`: `_wait("xz")` is
    called before x, z and mask are first read, `_wait("all")` before Y, v, h and s are, and
    `_done` as soon as a state is final.  With hooks the observation-noise statistics are taken
    after the state sequences (they do not depend on z, so the sweep is unchanged) — by then
    the large arrays have arrived."""
    import torch.distributed as td
    rank, world = _world(group)
    wait = _wait or (lambda g: None)
    done = _done or (lambda name, t: None)
    mark = _mark or (lambda label: None)
    defer_obs = _wait is not None
    st, pr = dict(states), dict(params)
    K = int(hyp["num_states"])
    k = data["Y"].shape[2]
    N, Tm = st["z"].shape
    wait("xz")
    F = G.ar_sufficient_stats(nlags=hyp["nlags"], num_states=K, mask=data["mask"], x=st["x"], z=st["z"])
    mark("ar_stats")
    counts = G.transition_counts(z=st["z"], mask=data["mask"], num_states=K)
    mark("counts")

    def obs_stats():
        return G.obs_variance_stats(Y=data["Y"], mask=data["mask"], Cd=pr["Cd"], v=st["v"], h=st["h"],
                                    x=st["x"], s=st["s"], Gamma=Gamma)
    S_y = msum = None
    if not defer_obs:
        S_y, msum = obs_stats()
        if world > 1:
            F, S_y, msum, counts = all_reduce_stats(F, S_y, msum, counts, group)
    elif world > 1:
        F, _, _, counts = all_reduce_stats(F, None, None, counts, group)
    # 1-3: replicated parameter draws (same key on every rank)
    pr["Ab"], pr["Q"] = G.sample_ar_params(key, F, nlags=hyp["nlags"], nu_0=hyp["nu_0"], S_0=hyp["S_0"],
                                           M_0=hyp["M_0"], K_0=hyp["K_0"])
    if not defer_obs:
        pr["sigmasq"] = G.sample_obs_variance(key, S_y, msum, nu_sigma=hyp["nu_sigma"], sigmasq_0=hyp["sigmasq_0"])
    n_glob = n_sessions_global if n_sessions_global is not None else N * world
    pr["betas"], pr["pi"] = G.sample_hdp_transitions(key, counts, pr["betas"], hyp["alpha"], hyp["kappa"],
                                                     hyp["gamma"], max_customers=n_glob * max(Tm - 1, 1))
    mark("params")
    # 4-8: per-session steps on the local shard; the Philox key is salted with the rank
    lkey = (key, "rank", rank)
    z_new = G.resample_stateseqs(lkey, x=st["x"], mask=data["mask"], Ab=pr["Ab"], Q=pr["Q"], pi=pr["pi"])[0]
    done("z", z_new)
    if defer_obs:
        wait("all")
        S_y, msum = obs_stats()     # uses the OLD x (st["x"]), exactly as step 2 of the sweep does
        if world > 1:
            S_y, msum = all_reduce_obs_stats(S_y, msum, group)
        pr["sigmasq"] = G.sample_obs_variance(key, S_y, msum, nu_sigma=hyp["nu_sigma"], sigmasq_0=hyp["sigmasq_0"])
    st["z"] = z_new
    st["x"] = G.resample_latents(lkey, Y=data["Y"], mask=data["mask"], v=st["v"], h=st["h"], z=st["z"],
                                 s=st["s"], Cd=pr["Cd"], sigmasq=pr["sigmasq"], Ab=pr["Ab"], Q=pr["Q"],
                                 Gamma=Gamma)
    done("x", st["x"])
    st["h"] = G.resample_heading(lkey, Y=data["Y"], v=st["v"], x=st["x"], s=st["s"], Cd=pr["Cd"],
                                 sigmasq=pr["sigmasq"], Gamma=Gamma)
    done("h", st["h"])
    st["v"] = G.resample_location(lkey, mask=data["mask"], Y=data["Y"], h=st["h"], x=st["x"], s=st["s"],
                                  Cd=pr["Cd"], sigmasq=pr["sigmasq"], sigmasq_loc=hyp["sigmasq_loc"],
                                  Gamma=Gamma)
    done("v", st["v"])
    st["s"] = G.resample_scales(lkey, x=st["x"], v=st["v"], h=st["h"], Y=data["Y"], Cd=pr["Cd"],
                                sigmasq=pr["sigmasq"], nu_s=hyp["nu_s"], s_0=hyp["s_0"], Gamma=Gamma)
    done("s", st["s"])
    return st, pr


_COPY_STREAMS = {}


def _copy_streams(dev):
    key = (dev.type, dev.index)
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
    return _COPY_STREAMS[key]


def gibbs_sweep_host(key, host_data, host_states, params, hyp, Gamma=None, group=None,
                     n_sessions_global=None, out=None, device=None, timeline=None):
    """One sweep driven from HOST arrays (the drop-in situation: the caller keeps data and states
    in host memory, e.g. as NumPy / JAX-CPU arrays).

    The uploads run on a copy stream in the order the sweep consumes them — x, z and mask first
    (sufficient statistics and the state sequences need nothing else), then Y, v, h and s, which
    arrive while those steps compute — and every new state is sent back on a second copy stream
    the moment it is final, overlapping the steps that follow.  Host tensors should be pinned;
    `out` (dict of pinned host tensors, allocated when None) receives z, x, h, v, s.
    Returns (out, params_on_device, states_on_device); the host copies are complete on return."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    main = torch.cuda.current_stream(dev)
    cs_in, cs_out = _copy_streams(dev)
    hd = {a: torch.as_tensor(b) for a, b in host_data.items()}
    hs = {a: torch.as_tensor(b) for a, b in host_states.items()}
    dd = {a: torch.empty(b.shape, dtype=b.dtype, device=dev) for a, b in hd.items()}
    ds = {a: torch.empty(b.shape, dtype=b.dtype, device=dev) for a, b in hs.items()}
    cs_in.wait_stream(main)     # the fresh buffers may alias memory that queued kernels still use
    first = [("s", "x"), ("s", "z"), ("d", "mask")]
    with torch.cuda.stream(cs_in):
        for kind, a in first:
            (ds if kind == "s" else dd)[a].copy_((hs if kind == "s" else hd)[a], non_blocking=True)
        ev_xz = torch.cuda.Event()
        ev_xz.record(cs_in)
        for a in hd:
            if ("d", a) not in first:
                dd[a].copy_(hd[a], non_blocking=True)
        for a in hs:
            if ("s", a) not in first:
                ds[a].copy_(hs[a], non_blocking=True)
        ev_all = torch.cuda.Event()
        ev_all.record(cs_in)
    if out is None:
        out = {}

    marks = []   # (label, event on the main stream) when a `timeline` dict is requested
    ev_start = torch.cuda.Event(enable_timing=True)
    if timeline is not None:
        ev_start.record(main)

    def mark(label):
        if timeline is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(main)
            marks.append((label, e))

    def wait(g):
        mark("before_wait_" + g)
        main.wait_event(ev_xz if (g == "xz" and os.environ.get("KPMS_E2E_NO_OVERLAP") != "1") else ev_all)
        mark("after_wait_" + g)

    def done(name, t):
        if name not in out:
            out[name] = torch.empty(t.shape, dtype=t.dtype).pin_memory()
        mark("done_" + name)
        ev = torch.cuda.Event()
        ev.record(main)
        cs_out.wait_event(ev)
        with torch.cuda.stream(cs_out):
            if out[name].dtype == t.dtype:
                out[name].copy_(t, non_blocking=True)
            else:
                out[name].copy_(t.to(out[name].dtype), non_blocking=True)
        t.record_stream(cs_out)

    st, pr = gibbs_sweep(key, dd, ds, params, hyp, Gamma=Gamma, group=group,
                         n_sessions_global=n_sessions_global, _wait=wait, _done=done, _mark=mark)
    cs_out.synchronize()
    if timeline is not None:
        main.synchronize()
        for label, e in marks:
            timeline[label] = ev_start.elapsed_time(e)
    return out, pr, st
