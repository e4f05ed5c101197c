"""Blocked Gibbs sweeps over session shards (one process per GPU).

Sessions are conditionally independent given the parameters, so every per-frame /
per-session step runs on its own shard with no communication.  The three parameter steps
consume only sufficient statistics of the previous sweep's state, which are packed into
one float64 buffer and combined with ONE all-reduce per sweep; every rank then draws
identical parameters from the same key (no broadcast).

`gibbs_sweep`  : the full keypoint-SLDS sweep, examples/keypoint_slds.ipynb cell 12 (steps 1-8):
                 ar_params, obs_variance, hdp_transitions, stateseqs, latents, heading, location, scales.
`arhmm_sweep`  : the AR-HMM sweep with x fixed, keypoint_slds.ipynb cell 9 / depth_ar_hmm.ipynb cell 8:
                 hdp_transitions, ar_params, stateseqs.

Random draws of session n depend on its GLOBAL index (`session_offset` + local index), never on
the rank, so z, x, h, v, s do not depend on how the sessions are sharded (the reference splits one
key per session: gibbs.py:20,54,112,145).

The chunk-parallel passes (HMM, Kalman) report a boundary mismatch.  A sweep launches all of its
steps optimistically and reads the mismatches back ONCE, at its end; a step that failed is redone
(with a longer warm-up, finally sequentially) together with the steps after it.  The warm-up
multiplier of the HMM passes is shared by all ranks: each rank's verdict on its last pass travels
in the packed all-reduce, so every rank changes it at the same sweep.
"""
import os

import torch

from . import gibbs as G

_N_FLAGS = 3   # [passes that failed, passes that were comfortable, passes reported] of the previous sweep


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


_OFFSETS = {}


def session_offset_of(n_local, group=None):
    """Global index of this rank's first session when shards are contiguous in rank order.
    Gathered once per (group, shard size) and cached: sharding is static during a fit."""
    import torch.distributed as td
    rank, world = _world(group)
    if world == 1:
        return 0
    key = (id(group), int(n_local))
    if key not in _OFFSETS:
        sizes = [None] * world
        td.all_gather_object(sizes, int(n_local), group=group)
        _OFFSETS[key] = sum(sizes[:rank])
    return _OFFSETS[key]


def pack_stats(F, S_y, msum, counts, flags=None):
    """-> flat float64 buffer [F | S_y | mask_sum | counts | flags]."""
    parts = [F.reshape(-1), S_y.reshape(-1), msum.reshape(-1), counts.to(torch.float64).reshape(-1)]
    if flags is not None:
        parts.append(flags.to(torch.float64).reshape(-1))
    return torch.cat(parts)


def unpack_stats(buf, K, P, k, n_flags=0):
    o = 0
    F = buf[o:o + K * P * P].reshape(K, P, P); o += K * P * P
    S_y = buf[o:o + k]; o += k
    msum = buf[o:o + 1]; o += 1
    counts = buf[o:o + K * K].reshape(K, K).round().to(torch.int64); o += K * K
    if n_flags:
        return F, S_y, msum, counts, buf[o:o + n_flags]
    return F, S_y, msum, counts


def all_reduce_stats(F, S_y, msum, counts, group=None, flags=None):
    """Sum the per-shard statistics over the ranks with ONE all-reduce of the packed buffer.
    S_y / msum may be None (they are then reduced later by `all_reduce_obs_stats`).  With `flags`
    (a small float64 vector) the summed flags are returned as a fifth value."""
    import torch.distributed as td
    K, P, k = counts.shape[0], F.shape[1], (S_y.numel() if S_y is not None else 1)
    buf = pack_stats(F, S_y if S_y is not None else F.new_zeros(k), msum if msum is not None else F.new_zeros(1),
                     counts, flags)
    td.all_reduce(buf, op=td.ReduceOp.SUM, group=group)
    out = unpack_stats(buf, K, P, k, 0 if flags is None else flags.numel())
    F, S2, m2, counts = out[:4]
    res = (F, (S2 if S_y is not None else None), (m2 if msum is not None else None), counts)
    return res if flags is None else res + (out[4],)


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


# ---------------------------------------------------------------------------------------------
# Shared warm-up multiplier of the chunked HMM passes.  `verdict` = what this rank saw in its last
# sweep (travels in the next all-reduce); `pending` = the summed verdicts of all ranks, a device
# tensor that is read back together with the boundary errors at the end of the sweep.
_HMM_SHARED = {"verdict": None}


def _local_verdict(dev):
    v = _HMM_SHARED["verdict"]
    if v is None:
        return torch.zeros(_N_FLAGS, dtype=torch.float64, device=dev)
    # built from Python scalars by fill kernels: no host->device copy that could wait for the stream
    return torch.stack([torch.full((), float(x), dtype=torch.float64, device=dev) for x in v])


def _apply_verdicts(failed, comfortable, reported):
    """Identical on every rank (the inputs are all-reduced sums)."""
    if reported <= 0:
        return
    G._hmm_adapt_update(failed > 0, comfortable >= reported)


def _verify_and_repair(checks, flags_dev, run_from, world, group=None, st=None):
    """ONE device->host read of everything a sweep left to verify; redo what failed.
    Sharded: the mismatches are also MAX-reduced over the ranks (a 3-float all-reduce enqueued before
    the read), because work drawn ahead for the next sweep (`next_key`) contains collectives: if ANY
    rank has to redo a step, EVERY rank discards what it drew ahead, so that all ranks issue the same
    collectives in the next sweep."""
    import torch.distributed as td
    dev = None
    for t in list(checks.values()) + [flags_dev]:
        if torch.is_tensor(t):
            dev = t.device
    if dev is None:
        if world == 1:
            return
        dev = torch.device("cuda", torch.cuda.current_device())
    # fixed layout [z mismatch, x filter mismatch, x sampler mismatch]; 0 where a pass ran sequentially
    zc = checks.get("z")
    xc = checks.get("x")
    local = torch.cat([zc.reshape(-1)[:1].to(torch.float64) if zc is not None else torch.zeros(1, dtype=torch.float64, device=dev),
                       xc.reshape(-1)[:2].to(torch.float64) if xc is not None else torch.zeros(2, dtype=torch.float64, device=dev)])
    local = torch.nan_to_num(local, nan=float("inf"), posinf=float("inf"))
    parts = [local]
    if world > 1:   # every rank, every sweep: the same collective whatever the local chunking decided
        glob = local.clone()
        td.all_reduce(glob, op=td.ReduceOp.MAX, group=group)
        parts.append(glob)
    if flags_dev is not None:
        parts.append(flags_dev.reshape(-1).to(torch.float64))
    host = torch.cat(parts).cpu().tolist()      # the sweep's only synchronisation
    vals = {}
    if zc is not None:
        vals["z"] = host[0]
    if xc is not None:
        vals["x"] = max(host[1], host[2])
    o = 3
    gvals = dict(vals)
    if world > 1:
        gvals = {"z": host[3], "x": max(host[4], host[5])}
        o = 6
    if flags_dev is not None:                    # verdicts of an earlier sweep, summed over the ranks
        _apply_verdicts(*host[o:o + _N_FLAGS])
    tolz, tolx = G.STATESEQ_CHUNKING["tol"], G.LATENTS_CHUNKING["tol"]
    if xc is not None or world > 1:
        # Kalman warm-up multiplier: moved by the GLOBAL mismatch, so every rank changes it at the same sweep
        G.note_latents_check(gvals["x"])
    any_failed = ("z" in gvals and not gvals["z"] <= tolz) or ("x" in gvals and not gvals["x"] <= tolx)
    if any_failed and st is not None:
        st.pop(NEXT_PARAMS_KEY, None)
        oc = st.get(OBS_STATS_KEY)
        if oc is not None and oc.get("reduced"):
            oc.update(S_y=oc["S_local"], msum=oc["m_local"], reduced=False)
    verdict = [0.0, 0.0, 0.0]
    z_failed = False
    if "z" in vals:
        b = vals["z"]
        verdict = [float(b > tolz), float(b <= 0.03 * tolz), 1.0]
        if b > tolz:
            z_failed = True
            G.FALLBACKS["stateseqs"] += 1
            G.FALLBACKS.setdefault("stateseqs_berr", []).append(b)
            del G.FALLBACKS["stateseqs_berr"][:-8]
    if world > 1:
        _HMM_SHARED["verdict"] = verdict
    else:
        _apply_verdicts(*verdict)
    def redo(A, step, max_boost):
        # the failed pass again, verified, with twice the warm-up (then the ladder of the step itself);
        # the multiplier proper moves through the verdicts only, identically on every rank
        saved = dict(A)
        A.update(boost=min(saved["boost"] * 2, max_boost), frozen=True)
        try:
            run_from(step, True)
        finally:
            A.clear()
            A.update(saved)

    if z_failed:
        redo(G._HMM_ADAPT, "z", G.HMM_MAX_BOOST)      # ... and everything that consumed the state sequences
    elif "x" in vals and not (vals["x"] <= tolx):
        G.FALLBACKS["latents"] += 1
        redo(G._LAT_ADAPT, "x", G.LATENTS_MAX_BOOST)


OBS_STATS_KEY = "_obs_stats"   # entry of the states dict that carries next sweep's (S_y, mask sum)


def _tensor_id(t):
    return (t.data_ptr(), t._version, tuple(t.shape)) if torch.is_tensor(t) else id(t)


def _obs_stats_tag(data, st, pr):
    """Identity of everything the observation-noise statistics depend on: they are reused only when
    the very tensors they were computed from come back unmodified."""
    return tuple(_tensor_id(t) for t in (data["Y"], data["mask"], st["x"], st["v"], st["h"], st["s"], pr["Cd"]))


NEXT_PARAMS_KEY = "_next_params"   # entry of the states dict: next sweep's Ab, Q, betas, pi, drawn ahead


def _draw_dynamics_params(key, data, st, pr, hyp, group, world, with_obs=None):
    """Steps 1 and 3 of a sweep from the states in `st`: AR statistics + transition counts (one packed
    all-reduce when sharded, which also carries `with_obs` = (S_y, mask sum) and the ranks' verdicts on
    their last chunked HMM pass), then the replicated draws of Ab, Q, betas, pi."""
    K = int(hyp["num_states"])
    F = G.ar_sufficient_stats(nlags=hyp["nlags"], num_states=K, mask=data["mask"], x=st["x"], z=st["z"])
    counts = G.transition_counts(z=st["z"], mask=data["mask"], num_states=K)
    S_y, msum = with_obs if with_obs is not None else (None, None)
    flags_dev = None
    if world > 1:
        F, S_y, msum, counts, flags_dev = all_reduce_stats(F, S_y, msum, counts, group, _local_verdict(F.device))
    out = {}
    out["Ab"], out["Q"] = G.sample_ar_params(key, F, nlags=hyp["nlags"], nu_0=hyp["nu_0"], S_0=hyp["S_0"],
                                             M_0=hyp["M_0"], K_0=hyp["K_0"])
    out["betas"], out["pi"] = G.sample_hdp_transitions(key, counts, pr["betas"], hyp["alpha"], hyp["kappa"],
                                                       hyp["gamma"])
    return out, S_y, msum, flags_dev


def _pose_steps_grouped(key, data, states, st, pr, hyp, Gamma, off, groups, checks, done, mark):
    """Steps 5-8 of a sweep (latents, heading, location, scales) one session group at a time: a group
    starts as soon as ITS data have arrived and its new states leave while the next group computes.
    Per-session steps with draws keyed by the global session index: the grouping changes nothing but
    the Kalman chunk layout (chosen per launch) and the summation order of the statistics."""
    parts = {a: [] for a in ("x", "h", "v", "s")}
    berrs, S_parts, m_parts = [], [], []
    for lo, hi, wait_group in groups:
        wait_group()
        sl = lambda t: t[lo:hi]
        kw = dict(Cd=pr["Cd"], sigmasq=pr["sigmasq"], Gamma=Gamma, session_offset=off + lo)
        x_g, info = G.resample_latents(key, Y=sl(data["Y"]), mask=sl(data["mask"]), v=sl(states["v"]),
                                       h=sl(states["h"]), z=sl(st["z"]), s=sl(states["s"]), Ab=pr["Ab"], Q=pr["Q"],
                                       strict="defer", return_info="check", **kw)
        if info["chunk_len"] > 0:
            berrs.append(info["boundary_err"].reshape(-1)[:2].to(torch.float64))
        done("x", x_g, lo, hi)
        h_g = G.resample_heading(key, Y=sl(data["Y"]), v=sl(states["v"]), x=x_g, s=sl(states["s"]), **kw)
        done("h", h_g, lo, hi)
        v_g = G.resample_location(key, mask=sl(data["mask"]), Y=sl(data["Y"]), h=h_g, x=x_g, s=sl(states["s"]),
                                  sigmasq_loc=hyp["sigmasq_loc"], **kw)
        done("v", v_g, lo, hi)
        s_g, (S_g, m_g) = G.resample_scales(key, x=x_g, v=v_g, h=h_g, Y=sl(data["Y"]), nu_s=hyp["nu_s"],
                                            s_0=hyp["s_0"], mask=sl(data["mask"]), return_stats=True, **kw)
        done("s", s_g, lo, hi)
        for a, t in (("x", x_g), ("h", h_g), ("v", v_g), ("s", s_g)):
            parts[a].append(t)
        S_parts.append(S_g)
        m_parts.append(m_g)
    mark("x")
    for a in parts:
        st[a] = torch.cat(parts[a], dim=0)
    checks["x"] = torch.stack(berrs).amax(0) if berrs else None
    S_loc, m_loc = torch.stack(S_parts).sum(0), torch.stack(m_parts).sum(0)
    st.pop(NEXT_PARAMS_KEY, None)
    st[OBS_STATS_KEY] = {"S_y": S_loc, "msum": m_loc, "reduced": False, "S_local": S_loc, "m_local": m_loc,
                         "tag": _obs_stats_tag(data, st, pr)}
    mark("hvs")


def gibbs_sweep(key, data, states, params, hyp, Gamma=None, group=None, n_sessions_global=None,
                session_offset=None, next_key=None, _wait=None, _done=None, _mark=None, _groups=None,
                _obs_carry=None):
    """One full sweep.  `data`/`states` hold this rank's shard; `params` are replicated.
    Returns new (states, params) dicts; inputs are not mutated.

    `session_offset`: global index of this shard's first session (default: gathered once from the
    shard sizes, contiguous in rank order).  `n_sessions_global` is accepted for compatibility.

    `next_key`: the key the NEXT sweep will be called with.  When given, that sweep's steps 1 and 3
    (AR statistics, transition counts, their all-reduce and the draws of Ab, Q, betas, pi — they depend
    on the new z and x only) are enqueued right after the latents, while the GPU still has the pose
    steps to run; the results travel in the returned states dict and are used by the next call if it
    presents the same tensors and the same key.  The sweep is unchanged; the launch latency of ~60 small
    kernels and the all-reduce leave the critical path.

    `_wait(group)` / `_done(name, tensor)` are the hooks of `gibbs_sweep_host`: `_wait("xz")` is
    called before x, z and mask are first read, `_wait("all")` before Y, v, h and s are, and
    `_done` as soon as a state is final.  With hooks the observation-noise statistics are taken
    after the state sequences (they do not depend on z, so the sweep is unchanged) — by then
    the large arrays have arrived.

    `_groups` / `_obs_carry` (host driver only): the per-session steps after the state sequences —
    latents, heading, location, scales — run session group by session group, `_groups` = [(lo, hi,
    wait)] with `wait()` blocking the stream until that group's Y, v, h, s have arrived, and every
    group's new states go back (`_done(name, tensor, lo, hi)`) while the next group computes.  That
    needs the observation-noise statistics before any of the data has arrived: `_obs_carry` =
    (S_y, mask sum), the statistics the previous sweep's scales pass left behind (already summed over
    the ranks).  Draws are keyed by the global session index, so grouping does not change them."""
    rank, world = _world(group)
    wait = _wait or (lambda g: None)
    done = _done or (lambda name, t, lo=None, hi=None: None)
    mark = _mark or (lambda label: None)
    defer_obs = _wait is not None
    st, pr = dict(states), dict(params)
    N = st["z"].shape[0]
    off = session_offset_of(N, group) if session_offset is None else int(session_offset)
    wait("xz")
    cached = st.pop(OBS_STATS_KEY, None)
    ahead = st.pop(NEXT_PARAMS_KEY, None)

    def obs_stats():
        # the scales step of the previous sweep has already accumulated these over its pass of Y
        if cached is not None and cached["tag"] == _obs_stats_tag(data, states, params):
            return cached["S_y"], cached["msum"], cached.get("reduced", False)
        S, m = G.obs_variance_stats(Y=data["Y"], mask=data["mask"], Cd=pr["Cd"], v=states["v"], h=states["h"],
                                    x=states["x"], s=states["s"], Gamma=Gamma)
        return S, m, False
    S_y = msum = None
    flags_dev = None
    if (ahead is not None and ahead["key"] == key and ahead["world"] == world and
            ahead["tag"] == tuple(_tensor_id(t) for t in (states["x"], states["z"], data["mask"], params["betas"]))):
        # steps 1 and 3 were drawn at the end of the previous sweep, from these very states (on a side stream)
        if ahead.get("event") is not None:
            cur = torch.cuda.current_stream(states["x"].device)
            cur.wait_event(ahead["event"])
            for t in list(ahead["params"].values()) + ([ahead["flags"]] if torch.is_tensor(ahead["flags"]) else []):
                t.record_stream(cur)
        pr.update(ahead["params"])
        flags_dev = ahead["flags"]
        if not defer_obs:
            S_y, msum, reduced = obs_stats()
            if world > 1 and not reduced:
                S_y, msum = all_reduce_obs_stats(S_y, msum, group)
    else:
        with_obs = None
        if not defer_obs:
            S_y, msum, reduced = obs_stats()
            if world > 1 and reduced:      # cannot be packed again: reduce the rest alone
                with_obs = None
            else:
                with_obs = (S_y, msum)
        drawn, S2, m2, flags_dev = _draw_dynamics_params(key, data, st, pr, hyp, group, world, with_obs)
        if with_obs is not None and world > 1:
            S_y, msum = S2, m2
        pr.update(drawn)
    mark("params_dyn")
    if not defer_obs:
        pr["sigmasq"] = G.sample_obs_variance(key, S_y, msum, nu_sigma=hyp["nu_sigma"], sigmasq_0=hyp["sigmasq_0"])
    mark("params")
    # 4-8: per-session steps on the local shard, launched without waiting for their verification
    x_old = st["x"]
    checks = {}

    def run_from(step, strict):
        if step == "z":
            z_new, _, info = G.resample_stateseqs(key, x=x_old, mask=data["mask"], Ab=pr["Ab"], Q=pr["Q"],
                                                  pi=pr["pi"], strict=strict, session_offset=off, return_info=True)
            checks["z"] = info["boundary_err"] if (strict == "defer" and info["chunk_len"] > 0) else None
            done("z", z_new)
            mark("z")
            if defer_obs and "sigmasq_done" not in checks:
                if _obs_carry is not None:
                    S_y, msum = _obs_carry            # left behind by the previous sweep, summed over the ranks
                else:
                    wait("all")
                    S_y, msum, reduced = obs_stats()     # uses the OLD x (st["x"]), exactly as step 2 of the sweep does
                    if world > 1 and not reduced:
                        S_y, msum = all_reduce_obs_stats(S_y, msum, group)
                pr["sigmasq"] = G.sample_obs_variance(key, S_y, msum, nu_sigma=hyp["nu_sigma"],
                                                      sigmasq_0=hyp["sigmasq_0"])
                checks["sigmasq_done"] = None
            st["z"] = z_new
        if _groups is not None and strict == "defer":
            _pose_steps_grouped(key, data, states, st, pr, hyp, Gamma, off, _groups, checks, done, mark)
            return
        wait("all")
        st["x"], info = G.resample_latents(key, Y=data["Y"], mask=data["mask"], v=states["v"], h=states["h"],
                                           z=st["z"], s=states["s"], Cd=pr["Cd"], sigmasq=pr["sigmasq"],
                                           Ab=pr["Ab"], Q=pr["Q"], Gamma=Gamma, strict=strict,
                                           session_offset=off, return_info="check")
        checks["x"] = info["boundary_err"] if (strict == "defer" and info["chunk_len"] > 0) else None
        done("x", st["x"])
        mark("x")
        st.pop(NEXT_PARAMS_KEY, None)
        if next_key is not None and not defer_obs and strict == "defer":
            # next sweep's steps 1 and 3 depend on the new z and x only: enqueue them now (not in a
            # repair pass, which runs on the failing ranks alone and must not issue collectives)
            # ... on a side stream: ~60 small launches (statistics, all-reduce, K-sized draws) that run next to
            # the frame-parallel pose kernels instead of between them
            main = torch.cuda.current_stream(st["x"].device)
            side = _side_stream(st["x"].device)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                nxt, _, _, nflags = _draw_dynamics_params(next_key, data, st, pr, hyp, group, world, None)
                ev_side = torch.cuda.Event()
                ev_side.record(side)
            for t in (st["x"], st["z"], data["mask"], pr["betas"]):
                t.record_stream(side)
            st[NEXT_PARAMS_KEY] = {"key": next_key, "world": world, "params": nxt, "flags": nflags, "event": ev_side,
                                   "tag": tuple(_tensor_id(t) for t in (st["x"], st["z"], data["mask"], pr["betas"]))}
            mark("ahead")
        st["h"] = G.resample_heading(key, Y=data["Y"], v=states["v"], x=st["x"], s=states["s"], Cd=pr["Cd"],
                                     sigmasq=pr["sigmasq"], Gamma=Gamma, session_offset=off)
        done("h", st["h"])
        st["v"] = G.resample_location(key, mask=data["mask"], Y=data["Y"], h=st["h"], x=st["x"], s=states["s"],
                                      Cd=pr["Cd"], sigmasq=pr["sigmasq"], sigmasq_loc=hyp["sigmasq_loc"],
                                      Gamma=Gamma, session_offset=off)
        done("v", st["v"])
        st["s"], (S_new, m_new) = G.resample_scales(key, x=st["x"], v=st["v"], h=st["h"], Y=data["Y"], Cd=pr["Cd"],
                                                    sigmasq=pr["sigmasq"], nu_s=hyp["nu_s"], s_0=hyp["s_0"],
                                                    Gamma=Gamma, session_offset=off, mask=data["mask"],
                                                    return_stats=True)
        reduced, S_loc, m_loc = False, S_new, m_new
        if world > 1 and next_key is not None and not defer_obs and strict == "defer":
            S_new, m_new = all_reduce_obs_stats(S_loc, m_loc, group)   # k + 1 doubles, off the next sweep's start
            reduced = True
        st[OBS_STATS_KEY] = {"S_y": S_new, "msum": m_new, "reduced": reduced, "S_local": S_loc, "m_local": m_loc,
                             "tag": _obs_stats_tag(data, st, pr)}
        done("s", st["s"])
        mark("hvs")

    run_from("z", "defer")
    _verify_and_repair(checks, flags_dev, run_from, world, group, st)
    mark("verified")
    return st, pr


def arhmm_sweep(key, data, states, params, hyp, group=None, session_offset=None):
    """One sweep of the AR-HMM with the continuous states fixed: resample_hdp_transitions,
    resample_ar_params, resample_stateseqs (keypoint_slds.ipynb cell 9; depth_ar_hmm.ipynb cell 8 —
    there `states['x']` is the data itself).  Sharded like `gibbs_sweep`: one packed all-reduce of
    the AR statistics and the transition counts."""
    rank, world = _world(group)
    st, pr = dict(states), dict(params)
    K = int(hyp["num_states"])
    N = st["z"].shape[0]
    off = session_offset_of(N, group) if session_offset is None else int(session_offset)
    F = G.ar_sufficient_stats(nlags=hyp["nlags"], num_states=K, mask=data["mask"], x=st["x"], z=st["z"])
    counts = G.transition_counts(z=st["z"], mask=data["mask"], num_states=K)
    flags_dev = None
    if world > 1:
        F, _, _, counts, flags_dev = all_reduce_stats(F, None, None, counts, group, _local_verdict(F.device))
    pr["betas"], pr["pi"] = G.sample_hdp_transitions(key, counts, pr["betas"], hyp["alpha"], hyp["kappa"],
                                                     hyp["gamma"])
    pr["Ab"], pr["Q"] = G.sample_ar_params(key, F, nlags=hyp["nlags"], nu_0=hyp["nu_0"], S_0=hyp["S_0"],
                                           M_0=hyp["M_0"], K_0=hyp["K_0"])
    checks = {}

    def run_from(step, strict):
        z_new, _, info = G.resample_stateseqs(key, x=st["x"], mask=data["mask"], Ab=pr["Ab"], Q=pr["Q"],
                                              pi=pr["pi"], strict=strict, session_offset=off, return_info=True)
        checks["z"] = info["boundary_err"] if (strict == "defer" and info["chunk_len"] > 0) else None
        st["z"] = z_new

    run_from("z", "defer")
    _verify_and_repair(checks, flags_dev, run_from, world, group, st)
    return st, pr


_COPY_STREAMS = {}
_SIDE_STREAMS = {}


def _side_stream(dev):
    key = (dev.type, dev.index)
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream(dev)
    return _SIDE_STREAMS[key]



def _copy_streams(dev):
    key = (dev.type, dev.index)
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
    return _COPY_STREAMS[key]


def _host_id(t):
    return (t.data_ptr(), t._version, tuple(t.shape))


def gibbs_sweep_host(key, host_data, host_states, params, hyp, Gamma=None, group=None,
                     n_sessions_global=None, session_offset=None, out=None, device=None, timeline=None,
                     resident=None, groups=None):
    """One sweep driven from HOST arrays (the drop-in situation: the caller keeps data and states
    in host memory, e.g. as NumPy / JAX-CPU arrays).

    The uploads run on a copy stream in the order the sweep consumes them — x, z and mask first
    (sufficient statistics and the state sequences need nothing else), then Y, v, h and s, which
    arrive while those steps compute — and every new state is sent back on a second copy stream
    the moment it is final, overlapping the steps that follow.  Host tensors should be pinned;
    `out` (dict of pinned host tensors, allocated when None) receives z, x, h, v, s.
    `resident`: optional dict of DEVICE tensors (e.g. {'Y': ..., 'mask': ...}) that the caller keeps
    on the GPU across sweeps; those arrays are not uploaded again.

    `groups` (default: 2 for one or two ranks, 4 when more ranks share the host — measured: one B200 65.3 /
    57.7 / 59.4 / 61.5 ms per sweep with 1 / 2 / 3 / 4 groups, eight B200 135 / 101 / 93 ms with 1 / 2 / 4):
    session-group pipelining.  When the states passed in are the very tensors the
    previous call wrote (the loop of a fit: `host_states = out` of the last sweep, untouched — checked
    by storage pointer and torch version counter, and agreed between the ranks), the observation-noise
    statistics of those states are already known from that call, so nothing in the sweep needs ALL of
    Y and s before it can start: the latents / heading / location / scales steps then run group by
    group, each as soon as its part of Y, v, h, s has arrived, and each group's results travel back
    while the next group computes.  Upload, compute and download overlap instead of following each
    other, which is what matters when several ranks share the host's PCIe bandwidth.  The first
    sweep of a loop (or any sweep whose states were modified) runs ungrouped.
    Returns (out, params_on_device, states_on_device); the host copies are complete on return."""
    import torch.distributed as td
    rank, world = _world(group)
    if groups is None:
        groups = 2 if world <= 2 else 4
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    main = torch.cuda.current_stream(dev)
    cs_in, cs_out = _copy_streams(dev)
    resident = resident or {}
    hd = {a: torch.as_tensor(b) for a, b in host_data.items() if a not in resident}
    hs = {a: torch.as_tensor(b) for a, b in host_states.items() if not a.startswith("_")}
    N = hs["x"].shape[0]
    # ---- may this sweep use what the previous one left behind?
    carry = host_states.get("_carry") if isinstance(host_states, dict) else None
    use = (groups > 1 and N >= 2 * groups and carry is not None and carry["world"] == world and
           all(_host_id(hs[a]) == carry["ids"].get(a) for a in ("x", "v", "h", "s")) and
           carry["Cd"] == _tensor_id(params["Cd"]) and
           carry["data"] == tuple((a, torch.as_tensor(b).data_ptr()) for a, b in sorted(host_data.items())))
    if world > 1:      # every rank takes the same path (the two paths issue different collectives)
        flag = torch.tensor([1.0 if use else 0.0], device=dev)
        td.all_reduce(flag, op=td.ReduceOp.MIN, group=group)
        use = bool(flag.item() > 0)
    bounds = [(g * N // groups, (g + 1) * N // groups) for g in range(groups)] if use else [(0, N)]

    dd = {a: torch.empty(b.shape, dtype=b.dtype, device=dev) for a, b in hd.items()}
    dd.update(resident)
    ds = {a: torch.empty(b.shape, dtype=b.dtype, device=dev) for a, b in hs.items()}
    cs_in.wait_stream(main)     # the fresh buffers may alias memory that queued kernels still use
    first = [("s", "x"), ("s", "z"), ("d", "mask")]
    ev_groups = []
    with torch.cuda.stream(cs_in):
        for kind, a in first:
            if kind == "d" and a not in hd:
                continue
            (ds if kind == "s" else dd)[a].copy_((hs if kind == "s" else hd)[a], non_blocking=True)
        ev_xz = torch.cuda.Event()
        ev_xz.record(cs_in)
        for lo, hi in bounds:                      # the rest, session group by session group
            for a in hd:
                if ("d", a) not in first:
                    dd[a][lo:hi].copy_(hd[a][lo:hi], non_blocking=True)
            for a in hs:
                if ("s", a) not in first:
                    ds[a][lo:hi].copy_(hs[a][lo:hi], non_blocking=True)
            e = torch.cuda.Event()
            e.record(cs_in)
            ev_groups.append(e)
        ev_all = ev_groups[-1]
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

    def done(name, t, lo=None, hi=None):
        if name not in out:
            out[name] = torch.empty((N,) + tuple(t.shape[1:]), dtype=t.dtype).pin_memory()
        mark("done_" + name)
        ev = torch.cuda.Event()
        ev.record(main)
        cs_out.wait_event(ev)
        dst = out[name] if lo is None else out[name][lo:hi]
        with torch.cuda.stream(cs_out):
            if dst.dtype == t.dtype:
                dst.copy_(t, non_blocking=True)
            else:
                dst.copy_(t.to(dst.dtype), non_blocking=True)
        t.record_stream(cs_out)

    grp = None
    if use:
        grp = [(lo, hi, (lambda e=e: main.wait_event(e))) for (lo, hi), e in zip(bounds, ev_groups)]
    st, pr = gibbs_sweep(key, dd, ds, params, hyp, Gamma=Gamma, group=group,
                         n_sessions_global=n_sessions_global, session_offset=session_offset,
                         _wait=wait, _done=done, _mark=mark, _groups=grp,
                         _obs_carry=(carry["S_y"], carry["msum"]) if use else None)
    # ---- what the next sweep may reuse: the statistics of the states just produced, summed over the ranks
    oc = st.get(OBS_STATS_KEY)
    new_carry = None
    if oc is not None:
        S_y, msum = oc["S_local"], oc["m_local"]
        if world > 1:
            S_y, msum = all_reduce_obs_stats(S_y, msum, group)
        new_carry = {"S_y": S_y, "msum": msum, "world": world, "Cd": _tensor_id(pr["Cd"]),
                     "data": tuple((a, torch.as_tensor(b).data_ptr()) for a, b in sorted(host_data.items()))}
    cs_out.synchronize()
    if new_carry is not None:
        new_carry["ids"] = {a: _host_id(out[a]) for a in ("x", "v", "h", "s")}
        out["_carry"] = new_carry
    else:
        out.pop("_carry", None)
    if timeline is not None:
        main.synchronize()
        for label, e in marks:
            timeline[label] = ev_start.elapsed_time(e)
    return out, pr, st


def arhmm_sweep_host(key, host_data, host_states, params, hyp, group=None, session_offset=None, out=None,
                     device=None):
    """`arhmm_sweep` driven from HOST arrays: x, z and mask go up (pinned buffers, copy stream), the
    new z comes back.  Returns (out, params_on_device, states_on_device)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    main = torch.cuda.current_stream(dev)
    cs_in, cs_out = _copy_streams(dev)
    cs_in.wait_stream(main)
    with torch.cuda.stream(cs_in):
        dd = {"mask": torch.as_tensor(host_data["mask"]).to(dev, non_blocking=True)}
        ds = {a: torch.as_tensor(host_states[a]).to(dev, non_blocking=True) for a in ("x", "z")}
    main.wait_stream(cs_in)
    for t in list(dd.values()) + list(ds.values()):
        t.record_stream(main)
    st, pr = arhmm_sweep(key, dd, ds, params, hyp, group=group, session_offset=session_offset)
    if out is None:
        out = {}
    if "z" not in out:
        out["z"] = torch.empty(st["z"].shape, dtype=st["z"].dtype).pin_memory()
    cs_out.wait_stream(main)
    with torch.cuda.stream(cs_out):
        out["z"].copy_(st["z"], non_blocking=True)
    st["z"].record_stream(cs_out)
    cs_out.synchronize()
    return out, pr, st
