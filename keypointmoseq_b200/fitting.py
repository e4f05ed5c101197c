"""A fit driver over the sweeps of `parallel.py`: the loops of examples/keypoint_slds.ipynb cell 9
(AR-only warm-up: resample_hdp_transitions, resample_ar_params, resample_stateseqs with x fixed)
and cell 12 (the full sweep), and of examples/depth_ar_hmm.ipynb cell 8 (the AR-HMM on given x),
with checkpoints.  The reference has no such function — its loops live in notebook cells — so the
names here are this package's own; the dict conventions (data / states / params / hyper-parameter
dicts) are the reference's.

Multi-GPU: call `fit` from every rank with that rank's shard of `data` / `states`; parameters are
replicated and stay identical (one packed all-reduce per sweep).  Rank 0 writes checkpoints of
parameters and history; every rank writes its own shard of the states.
"""
import os

import torch

from . import gibbs as G
from . import parallel


def _merge_hyp(*dicts, device=None):
    hyp = {}
    for d in dicts:
        for a, b in (d or {}).items():
            hyp[a] = torch.as_tensor(b, device=device) if hasattr(b, "shape") and getattr(b, "ndim", 0) > 0 else b
    return hyp


def to_device(d, device, int_keys=("z",)):
    """dict of arrays -> dict of device tensors in the storage dtypes of the kernels."""
    out = {}
    for a, b in d.items():
        t = torch.as_tensor(b)
        if a in int_keys:
            out[a] = t.to(device=device, dtype=torch.int32)
        elif a in ("Ab", "Q", "pi", "betas", "Cd", "sigmasq"):
            out[a] = t.to(device=device, dtype=torch.float64)
        else:
            out[a] = t.to(device=device, dtype=torch.float32)
    return out


def checkpoint_path(directory, rank=0):
    return os.path.join(directory, "checkpoint_rank%d.pt" % rank)


def save_checkpoint(directory, iteration, states, params, history, rank=0, phase="full"):
    """torch.save of the three dicts (the reference's state is exactly these dicts)."""
    os.makedirs(directory, exist_ok=True)
    path = checkpoint_path(directory, rank)
    tmp = path + ".tmp"
    torch.save({"iteration": int(iteration), "phase": phase,
                "states": {a: b.detach().cpu() for a, b in states.items() if torch.is_tensor(b)},
                "params": {a: b.detach().cpu() for a, b in params.items()},
                "history": history,
                # warm-up multipliers of the chunked passes: a resumed fit continues exactly where it stopped
                "adapt": {"hmm": dict(G._HMM_ADAPT), "kalman": dict(G._LAT_ADAPT)}}, tmp)
    os.replace(tmp, path)     # a crash mid-write leaves the previous checkpoint intact
    return path


def load_checkpoint(directory, rank=0, device="cuda"):
    path = checkpoint_path(directory, rank)
    if not os.path.exists(path):
        return None
    ck = torch.load(path, map_location="cpu", weights_only=False)
    ck["states"] = to_device(ck["states"], device)
    ck["params"] = to_device(ck["params"], device)
    if "adapt" in ck:
        G._HMM_ADAPT.update(ck["adapt"]["hmm"])
        G._LAT_ADAPT.update(ck["adapt"]["kalman"])
    return ck


def fit(key, data, states, params, *, trans_hypparams, ar_hypparams, obs_hypparams=None,
        translation_hypparams=None, num_iters=50, ar_only_iters=0, ar_only_trans_hypparams=None,
        Gamma=None, group=None, session_offset=None, checkpoint_dir=None, checkpoint_every=25,
        resume=True, log_every=0, log_fn=print):
    """Run `ar_only_iters` AR-HMM sweeps (x fixed) followed by `num_iters` sweeps.

    With `obs_hypparams` given the second phase is the full keypoint-SLDS sweep (data needs 'Y' and
    'mask', states need x, z, v, h, s); without it every sweep is the AR-HMM sweep on
    `states['x']` (the depth_ar_hmm.ipynb model).  `ar_only_trans_hypparams` are the transition
    hyper-parameters of the warm-up phase (keypoint_slds.ipynb cell 4 vs cell 11); default: the same.
    Sweep i uses the key (key, i), in both phases, as the notebooks reuse `keys[i]`.

    Checkpoints (`checkpoint_dir`): every `checkpoint_every` sweeps and at the end; with
    `resume=True` an existing checkpoint is loaded and the loop continues after its iteration.
    Returns (states, params, history); history holds per-logged-sweep diagnostics."""
    rank, world = parallel._world(group)
    dev = states["x"].device
    full = obs_hypparams is not None
    hyp_full = _merge_hyp(trans_hypparams, ar_hypparams, obs_hypparams, translation_hypparams, device=dev)
    hyp_ar = _merge_hyp(ar_only_trans_hypparams or trans_hypparams, ar_hypparams, device=dev)
    total = int(ar_only_iters) + int(num_iters)
    st, pr = dict(states), dict(params)
    history, start = [], 0
    if checkpoint_dir and resume:
        ck = load_checkpoint(checkpoint_dir, rank, dev)
        if ck is not None:
            st, pr, history, start = ck["states"], ck["params"], ck["history"], ck["iteration"] + 1
    if session_offset is None:
        session_offset = parallel.session_offset_of(st["z"].shape[0], group)
    for i in range(start, total):
        warm = i < ar_only_iters
        if warm or not full:
            st, pr = parallel.arhmm_sweep((key, i), data, st, pr, hyp_ar if warm else hyp_full, group=group,
                                          session_offset=session_offset)
        else:
            st, pr = parallel.gibbs_sweep((key, i), data, st, pr, hyp_full, Gamma=Gamma, group=group,
                                          session_offset=session_offset,
                                          next_key=(key, i + 1) if i + 1 < total else None)
        if log_every and (i % log_every == 0 or i == total - 1):
            usage, durations = G.stateseq_stats(st["z"], data["mask"])
            rec = {"iteration": i, "phase": "ar_only" if warm else ("full" if full else "arhmm"),
                   "states_used": int((usage > 0).sum()),
                   "median_duration": float(durations.to(torch.float64).median()) if durations.numel() else 0.0}
            history.append(rec)
            if rank == 0 and log_fn is not None:
                log_fn(rec)
        if checkpoint_dir and ((i + 1) % int(checkpoint_every) == 0 or i == total - 1):
            save_checkpoint(checkpoint_dir, i, st, pr, history, rank, "ar_only" if warm else "full")
    st.pop(parallel.OBS_STATS_KEY, None)   # the sweeps' private caches of next sweep's statistics / parameters
    st.pop(parallel.NEXT_PARAMS_KEY, None)
    return st, pr, history
