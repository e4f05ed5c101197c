"""Blocked Gibbs steps of the keypoint-SLDS — same names, keyword arguments and return
values as the reference's `keypoint_moseq/gibbs.py`, on torch CUDA tensors, calling the
sm_100a kernels of libkpms_b200.so through the C-ABI (include/kpms_b200.h).

Conventions kept from the reference: `resample_X(key, *, <named arrays>, **kwargs)`;
unknown keyword arguments are swallowed so callers can splat the data / states / params /
hyper-parameter dicts; inputs are never mutated; outputs are fresh tensors.

Differences: per-frame arrays are stored float32 (parameters and
statistics float64, state sequences int32); `key` is an int / tuple / torch.Generator;
every step takes an optional `draws={...}` dict of pre-generated variates (parity tests
inject the oracle's draws) and an optional `Gamma=` centering basis.  There is no CPU
fallback: a missing library or a CPU tensor raises.
"""
import torch

from . import _lib
from . import distributions as dist
from .util import center_embedding, expand_Cd, f32, f64, i32

# Kalman time-chunking: "auto" cuts every session into chunks sized so that one wave of warps
# covers the shard; each chunk is warmed up over `burn_in` earlier steps.  With strict=True the
# reported chunk-boundary mismatch is read back and the call is repeated in the exact sequential
# mode (chunk_len = 0) if it exceeds `tol` (never observed at burn_in >= 128).
LATENTS_CHUNKING = {"chunk_len": "auto", "burn_in": 256, "strict": True, "tol": 1e-10}
KALMAN_WARPS_PER_SM = 12
last_latents_boundary_err = None
# HMM chunking (K <= 128): same idea for the forward filter and the backward sampler.  The library
# reports the largest absolute mismatch of the normalised messages at chunk boundaries (1e30 for a
# mask that is not trailing padding); sampler chunks are repaired on the device and never fail.
# `tol`: a mismatch eps perturbs only the first frames of a chunk and decays geometrically.
# Measured on the bench shard (tools/hmm_model_data.py, 17.5 k chunks): 1e-4 changes neither the
# sampled states nor the marginal log-likelihood; 5e-2 moves the log-likelihood by 6e-6 relative
# (the contract allows 1e-5) and 3e-5 of the states (contract: 1e-3).  1e-3 keeps a factor 50 of
# margin; above it strict mode retries with a longer warm-up, then runs the sequential kernels.
STATESEQ_CHUNKING = {"chunk_len": "auto", "burn_in": 256, "strict": True, "tol": 1e-3}
HMM_CTAS_PER_SM = 3
last_stateseq_boundary_err = None
# how often a chunked pass failed its boundary check and was redone exactly (diagnostics)
FALLBACKS = {"stateseqs": 0, "latents": 0}


def auto_hmm_chunk_len(N, Tm, K, burn_in, device):
    """Chunk length of the forward filter.  K <= 104 runs on the tensor cores (states padded to a multiple of 4),
    128 chains per CTA: aim at one CTA per SM but keep a chunk at least as long as its warm-up.
    Other K use the CUDA-core kernel (one CTA per chain, three resident per SM), where longer
    chunks keep the warm-up overhead small."""
    sms = torch.cuda.get_device_properties(device).multi_processor_count
    if K > 512:
        return 0                      # sequential K-generic kernel
    if K > 128:                       # 32 chains per CTA (hmm_big.cu): one CTA per SM, chunks no shorter than their warm-up
        nch = max(1, sms * 32 // max(N, 1))
        cl = max(-(-Tm // nch), max(burn_in, 1))
    elif K <= 104:
        nch = max(1, sms * 128 // max(N, 1))
        cl = max(-(-Tm // nch), max(burn_in, 1))
    else:
        nch = max(1, sms * 3 // max(N, 1))
        cl = max(-(-Tm // nch), 4 * max(burn_in, 1))
    return 0 if 2 * cl >= Tm else cl


def auto_chunk_len(N, T, L, burn_in, device):
    """Kalman chunk length such that N * n_chunks is just under one wave of resident warps.  With few
    sessions the wave cannot be filled with long chunks: the chain latency (chunk + warm-up steps) is
    then what counts, and chunks shrink down to twice the warm-up."""
    Tt = T - L
    slots = torch.cuda.get_device_properties(device).multi_processor_count * KALMAN_WARPS_PER_SM
    nch = max(1, slots // max(N, 1))
    cl = max(-(-Tt // nch), 2 * max(burn_in, 1), 256)
    return 0 if cl >= Tt else cl


def latents_check_ok(berr):
    """Reads the Kalman chunk-boundary mismatch back (synchronises); NaN counts as a failure."""
    b = float(berr.max()) if torch.is_tensor(berr) else float(berr)
    return b == b and b <= LATENTS_CHUNKING["tol"]


def spd_inv(A):
    """Inverse of symmetric positive definite matrices through their Cholesky factor.  (torch.linalg.inv
    goes through MAGMA's batched LU for small batches, which synchronises the whole DEVICE and so
    stalls behind copies running on other streams; potrf / trsm stay on the current stream.
    cholesky_ex does not read the status back, so nothing here synchronises with the host: like the
    reference, a non-PD input yields NaNs rather than an exception.)"""
    L = torch.linalg.cholesky_ex(A)[0]
    eye = torch.eye(A.shape[-1], dtype=A.dtype, device=A.device).expand(A.shape)
    Li = torch.linalg.solve_triangular(L, eye, upper=False)
    return Li.transpose(-1, -2) @ Li


def _dims(Ab):
    K, D = Ab.shape[0], Ab.shape[1]
    L = Ab.shape[2] // Ab.shape[1]          # gibbs.py:12,102,109
    return K, D, L


def _gamma(Gamma, k, dev):
    return center_embedding(k, dev) if Gamma is None else f64(Gamma, dev)


# ------------------------------------------------------------------------ a1-a4
def ar_log_likelihood(*, x, Ab, Q, impl=0, **kwargs):
    """gibbs.py:100-104,110 -> ll (N, T-L, K) float32."""
    lib = _lib.load()
    x = f32(x)
    dev = x.device
    Ab, Q = f64(Ab, dev), f64(Q, dev)
    K, D, L = _dims(Ab)
    N, T, _ = x.shape
    ll = torch.empty((N, T - L, K), dtype=torch.float32, device=dev)
    nbytes = lib.kpms_ar_loglik_workspace_bytes(K, D, L)
    ws = _lib.WORKSPACE.get("ar_loglik", nbytes, dev)
    _lib.check(lib.kpms_ar_loglik(_lib.ptr(x), _lib.ptr(Ab), _lib.ptr(Q), N, T, D, L, K, _lib.ptr(ll),
                                  _lib.ptr(ws), ws.numel(), int(impl), _lib.stream_ptr(dev)),
               "kpms_ar_loglik")
    return ll


# Adaptive warm-up of the chunked HMM passes: how fast the filter forgets depends on the current
# parameters (slowly early in a fit, when the states are hardly distinguishable).  A failed boundary
# check doubles the warm-up for this and the following calls; a run of comfortable passes halves it
# again.  `boost` multiplies the configured burn_in.
_HMM_ADAPT = {"boost": 1, "streak": 0, "need": 8, "probing": False}


SPORADIC_GAP = 6


def _adapt_update(A, failed, comfortable, max_boost, min_boost=1):
    """One verdict on a chunked pass -> the warm-up multiplier of the next passes.
    A failed pass is always redone with twice the warm-up (by the caller).  The multiplier itself
    doubles only when failures come close together (two within SPORADIC_GAP passes) or when a probe
    of a shorter warm-up fails: a lone failure — one chunk boundary among thousands that needed
    longer, as happens every few sweeps with K = 500 nearly identical states — costs its repeat and
    nothing more.  After `need` consecutive comfortable passes (mismatch < 3 % of the tolerance) the
    multiplier is halved again, on probation: if the shorter warm-up then fails, `need` is quadrupled
    (cap 4096), so a model that needs the longer warm-up is re-probed ever more rarely."""
    if A.get("frozen"):
        return
    if failed:
        recent = A.get("since_fail", 10 ** 9) < SPORADIC_GAP
        A["since_fail"] = 0
        if A["probing"]:
            A["need"] = min(A["need"] * 4, 4096)
        if A["probing"] or recent:
            dbl = A["boost"] * 2
            A["boost"] = min(int(dbl) if dbl >= 1 else dbl, max_boost)
        A["streak"], A["probing"] = 0, False
        return
    A["since_fail"] = A.get("since_fail", 10 ** 9) + 1
    if comfortable:
        A["streak"] += 1
        if A["probing"] and A["streak"] >= 8:
            A["probing"] = False
        if A["streak"] >= A["need"] and A["boost"] > min_boost:
            half = A["boost"] / 2
            A["boost"] = int(half) if half >= 1 else half
            A["streak"], A["probing"] = 0, True
    else:
        A["streak"] = 0


def _hmm_adapt_update(failed, comfortable):
    _adapt_update(_HMM_ADAPT, failed, comfortable, HMM_MAX_BOOST)


# the same book-keeping for the chunked Kalman filter / sampler (multiplies LATENTS_CHUNKING["burn_in"] when
# the caller does not pass a warm-up of its own)
# The Kalman multiplier may also go BELOW one: the default warm-up (256 steps) is safe for every model
# measured (mismatch at rounding level), half of it is enough where the observations are informative
# (16 keypoints: 4e-14 at 128 steps) but not where they are few (8 keypoints: above the tolerance), so a
# fit starts safe and tries the shorter warm-up after `need` comfortable sweeps.
_LAT_ADAPT = {"boost": 1, "streak": 0, "need": 3, "probing": False}
LATENTS_MAX_BOOST = 16
LATENTS_MIN_BOOST = 0.5


def note_latents_check(b):
    """Book-keeping after a chunked Kalman pass with boundary mismatch `b` (NaN = failure)."""
    tol = LATENTS_CHUNKING["tol"]
    ok = b == b and b <= tol
    _adapt_update(_LAT_ADAPT, not ok, ok and b <= 0.03 * tol, LATENTS_MAX_BOOST, LATENTS_MIN_BOOST)
    return ok
HMM_MAX_BOOST = 16


def _hmm_burn(burn_in, boost):
    return burn_in if boost == 1 else max(burn_in, 32) * boost


def note_hmm_check(b, adaptive=True):
    """Book-keeping of the adaptive warm-up after a chunked pass whose boundary mismatch was `b`.
    Returns True when the pass is acceptable."""
    tol = STATESEQ_CHUNKING["tol"]
    if b <= tol:
        if adaptive:
            _hmm_adapt_update(False, b <= 0.03 * tol)
        return True
    FALLBACKS["stateseqs"] += 1
    FALLBACKS.setdefault("stateseqs_berr", []).append(b)
    del FALLBACKS["stateseqs_berr"][:-8]
    return False


def _run_hmm_adaptive(run, berr, chunk_len, burn_in, strict, adaptive):
    """run(chunk_len, burn_in) launches one chunked pass and fills `berr`.
    strict=True reads the mismatch back (one host synchronisation) and repeats the pass with a longer
    warm-up, finally with the sequential kernels, until it is within tolerance.  strict="defer"
    launches once with the current adaptive warm-up and leaves the check to the caller
    (`parallel.gibbs_sweep` reads all boundary errors of a sweep back together, once)."""
    if chunk_len <= 0 or not strict:
        run(chunk_len, burn_in if chunk_len > 0 else 0)
        return
    boost = _HMM_ADAPT["boost"] if adaptive else 1
    if strict == "defer":
        run(chunk_len, _hmm_burn(burn_in, boost))
        return
    while True:
        run(chunk_len, _hmm_burn(burn_in, boost))
        b = float(berr)
        if b != b:
            b = float("inf")
        if note_hmm_check(b, adaptive):
            return
        boost *= 2
        if b > 1e29 or boost > HMM_MAX_BOOST:     # 1e30 flags a mask that is not trailing padding
            FALLBACKS["stateseqs_exact"] = FALLBACKS.get("stateseqs_exact", 0) + 1
            run(0, 0)
            return
        if adaptive:
            _hmm_adapt_update(True, False)
            boost = max(boost, _HMM_ADAPT["boost"])


def _hmm_chunking(N, Tm, K, chunk_len, burn_in, strict, dev):
    chunk_len = STATESEQ_CHUNKING["chunk_len"] if chunk_len is None else chunk_len
    burn_in = STATESEQ_CHUNKING["burn_in"] if burn_in is None else burn_in
    strict = STATESEQ_CHUNKING["strict"] if strict is None else strict
    auto = chunk_len == "auto"
    if auto:
        chunk_len = auto_hmm_chunk_len(N, Tm, K, int(burn_in), dev)
    return int(chunk_len), int(burn_in), (strict if strict == "defer" else bool(strict)), auto


def sample_hmm_stateseqs(*, ll, mask, pi, u, chunk_len=None, burn_in=None, strict=None, return_info=False):
    """distributions.py:46-79 batched over sessions: ll (N,Tm,K), mask (N,Tm), u (N,Tm)."""
    global last_stateseq_boundary_err
    lib = _lib.load()
    ll = f32(ll)
    dev = ll.device
    N, Tm, K = ll.shape
    mask, pi, u = f32(mask, dev), f64(pi, dev), f32(u, dev)
    chunk_len, burn_in, strict, auto = _hmm_chunking(N, Tm, K, chunk_len, burn_in, strict, dev)
    z = torch.empty((N, Tm), dtype=torch.int32, device=dev)
    loglik = torch.empty((N,), dtype=torch.float64, device=dev)
    berr = torch.zeros((1,), dtype=torch.float32, device=dev)

    def run(cl, bi):
        ws = _lib.WORKSPACE.get("hmm", lib.kpms_hmm_ffbs_workspace_bytes(N, Tm, K, cl), dev)
        _lib.check(lib.kpms_hmm_ffbs(_lib.ptr(ll), _lib.ptr(mask), Tm, _lib.ptr(pi), _lib.ptr(u), N, Tm, K, cl,
                                     bi, _lib.ptr(z), _lib.ptr(loglik), _lib.ptr(berr), _lib.ptr(ws),
                                     ws.numel(), _lib.stream_ptr(dev)), "kpms_hmm_ffbs")

    _run_hmm_adaptive(run, berr, chunk_len, burn_in, strict, auto)
    last_stateseq_boundary_err = berr
    if return_info:
        return z, loglik, {"boundary_err": berr}
    return z, loglik


def resample_stateseqs(key, *, x, mask, Ab, Q, pi, draws=None, impl=0, chunk_len=None, burn_in=None,
                       strict=None, session_offset=0, return_info=False, **kwargs):
    """gibbs.py:107-115 -> (z (N,T-L) int32, marginal log-likelihood (N,) float64).
    `session_offset`: global index of this shard's first session (keys the Philox draws, so that
    the samples do not depend on the sharding)."""
    global last_stateseq_boundary_err
    lib = _lib.load()
    x = f32(x)
    dev = x.device
    mask, Ab, Q, pi = f32(mask, dev), f64(Ab, dev), f64(Q, dev), f64(pi, dev)
    K, D, L = _dims(Ab)
    N, T, _ = x.shape
    if draws is not None:
        u = f32(draws["u"], dev)
    else:
        u = dist.fill_uniform((N, T - L), dist.seed_of(key), dist.STREAM_Z, dev, session_offset)
    chunk_len, burn_in, strict, auto = _hmm_chunking(N, T - L, K, chunk_len, burn_in, strict, dev)
    z = torch.empty((N, T - L), dtype=torch.int32, device=dev)
    loglik = torch.empty((N,), dtype=torch.float64, device=dev)
    berr = torch.zeros((1,), dtype=torch.float32, device=dev)

    Tm = T - L

    def run(cl, bi):
        ws = _lib.WORKSPACE.get("stateseqs", lib.kpms_resample_stateseqs_workspace_bytes(N, T, D, L, K, cl), dev)
        _lib.check(lib.kpms_resample_stateseqs(
            _lib.ptr(x), _lib.ptr(mask), _lib.ptr(Ab), _lib.ptr(Q), _lib.ptr(pi), _lib.ptr(u), N, T, D, L, K,
            cl, bi, _lib.ptr(z), _lib.ptr(loglik), _lib.ptr(berr), _lib.ptr(ws), ws.numel(), int(impl),
            _lib.stream_ptr(dev)), "kpms_resample_stateseqs")

    _run_hmm_adaptive(run, berr, chunk_len, burn_in, strict, auto)
    last_stateseq_boundary_err = berr
    if return_info:
        return z, loglik, {"boundary_err": berr, "chunk_len": chunk_len}
    return z, loglik


# ------------------------------------------------------------------------ a5-a7
def resample_latents(key, *, Y, mask, v, h, z, s, Cd, sigmasq, Ab, Q, Gamma=None, draws=None,
                     chunk_len=None, burn_in=None, strict=None, return_info=False, session_offset=0,
                     **kwargs):
    """gibbs.py:10-23 -> x (N,T,D) float32.  With `return_info=True` also returns a dict with
    the innovation-form marginal log-likelihood (N,) and the chunk-boundary mismatch (2,)."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    mask, v, h, s = f32(mask, dev), f32(v, dev), f32(h, dev), f32(s, dev)
    z = i32(z, dev)
    Ab, Q, sigmasq = f64(Ab, dev), f64(Q, dev), f64(sigmasq, dev)
    K, D, L = _dims(Ab)
    n = D * L
    Cdfull = expand_Cd(f64(Cd, dev), _gamma(Gamma, k, dev), dk)
    if draws is not None:
        w, w_last = f32(draws["w"], dev), f32(draws["w_last"], dev)
    else:
        seed = dist.seed_of(key)
        w = dist.fill_normal((N, T - L, n), seed, dist.STREAM_X, dev, session_offset)
        w_last = dist.fill_normal((N, n), seed, dist.STREAM_X_LAST, dev, session_offset)
    global last_latents_boundary_err
    chunk_len = LATENTS_CHUNKING["chunk_len"] if chunk_len is None else chunk_len
    adaptive = burn_in is None       # default warm-up: lengthened for models that were seen to forget slowly
    burn_in = int(LATENTS_CHUNKING["burn_in"] * _LAT_ADAPT["boost"]) if adaptive else burn_in
    strict = LATENTS_CHUNKING["strict"] if strict is None else strict
    if chunk_len == "auto":
        chunk_len = auto_chunk_len(N, T, L, int(burn_in), dev)
    x = torch.empty((N, T, D), dtype=torch.float32, device=dev)
    # return_info="check": the boundary mismatch only (a sweep has no use for the marginal log-likelihood,
    # which costs the filter a product over the pivots and a float64 log per step)
    loglik = torch.empty((N,), dtype=torch.float64, device=dev) if return_info is True else None
    berr = torch.zeros((2,), dtype=torch.float64, device=dev)

    def run(cl, bi):
        nbytes = lib.kpms_resample_latents_workspace_bytes(N, T, k, dk, D, L, K, int(cl))
        ws = _lib.WORKSPACE.get("latents", nbytes, dev)
        _lib.check(lib.kpms_resample_latents(
            _lib.ptr(Y), _lib.ptr(mask), _lib.ptr(v), _lib.ptr(h), _lib.ptr(z), _lib.ptr(s), _lib.ptr(Cdfull),
            _lib.ptr(sigmasq), _lib.ptr(Ab), _lib.ptr(Q), _lib.ptr(w), _lib.ptr(w_last), N, T, k, dk, D, L, K,
            int(cl), int(bi), _lib.ptr(x), _lib.ptr(loglik), _lib.ptr(berr), _lib.ptr(ws),
            ws.numel(), _lib.stream_ptr(dev)), "kpms_resample_latents")

    run(chunk_len, burn_in)
    if chunk_len > 0 and strict and strict != "defer" and not (
            note_latents_check(float(berr.max())) if adaptive else latents_check_ok(berr)):
        # warm-up too short for this model: once more with four times the warm-up, then exactly
        # (one warp per session)
        FALLBACKS["latents"] += 1
        burn2 = 4 * max(int(burn_in), 64)
        cl2 = auto_chunk_len(N, T, L, burn2, dev)
        ok = False
        if cl2 > 0:
            run(cl2, burn2)
            ok = latents_check_ok(berr)
        if not ok:
            FALLBACKS["latents_exact"] = FALLBACKS.get("latents_exact", 0) + 1
            run(0, 0)
    last_latents_boundary_err = berr
    if return_info:
        return x, {"loglik": loglik, "boundary_err": berr, "chunk_len": chunk_len}
    return x


# ------------------------------------------------------------------------ a15
def resample_heading(key, *, Y, v, x, s, Cd, sigmasq, Gamma=None, draws=None, return_info=False,
                     session_offset=0, **kwargs):
    """gibbs.py:25-35 -> h (N,T) float32 in [-pi, pi]."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    v, x, s, sigmasq = f32(v, dev), f32(x, dev), f32(s, dev), f64(sigmasq, dev)
    D = x.shape[-1]
    Cdfull = expand_Cd(f64(Cd, dev), _gamma(Gamma, k, dev), dk)
    hout = torch.empty((N, T), dtype=torch.float32, device=dev)
    theta = torch.empty((N, T), dtype=torch.float64, device=dev) if return_info else None
    kappa = torch.empty((N, T), dtype=torch.float64, device=dev) if return_info else None
    info = {"theta": theta, "kappa": kappa}
    if draws is not None:
        # injected proposals: u (N,T,rounds,3); a frame whose rounds are all rejected gets the mean direction
        u = f32(draws["u"], dev)
        rounds = u.shape[2]
        _lib.check(lib.kpms_resample_heading(
            _lib.ptr(Y), _lib.ptr(v), _lib.ptr(x), _lib.ptr(s), _lib.ptr(Cdfull), _lib.ptr(sigmasq), _lib.ptr(u),
            rounds, N, T, k, dk, D, _lib.ptr(hout), _lib.ptr(theta), _lib.ptr(kappa), _lib.stream_ptr(dev)),
            "kpms_resample_heading")
    else:
        # uniforms drawn inside the kernel, keyed by the global frame index; the rejection loop runs
        # until it accepts (no pre-drawn round limit)
        exhausted = torch.zeros((1,), dtype=torch.int32, device=dev) if return_info else None
        _lib.check(lib.kpms_resample_heading_philox(
            _lib.ptr(Y), _lib.ptr(v), _lib.ptr(x), _lib.ptr(s), _lib.ptr(Cdfull), _lib.ptr(sigmasq),
            dist.seed_of(key), dist.STREAM_H, int(session_offset) * T, N, T, k, dk, D, _lib.ptr(hout),
            _lib.ptr(theta), _lib.ptr(kappa), _lib.ptr(exhausted), _lib.stream_ptr(dev)),
            "kpms_resample_heading_philox")
        info["exhausted"] = exhausted
    if return_info:
        return hout, info
    return hout


# ------------------------------------------------------------------------ a16
def resample_location(key, *, mask, Y, h, x, s, Cd, sigmasq, sigmasq_loc, Gamma=None, draws=None,
                      return_info=False, session_offset=0, **kwargs):
    """gibbs.py:38-54 -> v (N,T,dk) float32."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    mask, h, x, s, sigmasq = f32(mask, dev), f32(h, dev), f32(x, dev), f32(s, dev), f64(sigmasq, dev)
    D = x.shape[-1]
    Cdfull = expand_Cd(f64(Cd, dev), _gamma(Gamma, k, dev), dk)
    if draws is not None:
        w, w_last = f32(draws["w"], dev), f32(draws["w_last"], dev)
    else:
        seed = dist.seed_of(key)
        w = dist.fill_normal((N, T - 1, dk), seed, dist.STREAM_V, dev, session_offset)
        w_last = dist.fill_normal((N, dk), seed, dist.STREAM_V_LAST, dev, session_offset)
    vout = torch.empty((N, T, dk), dtype=torch.float32, device=dev)
    loglik = torch.empty((N,), dtype=torch.float64, device=dev) if return_info else None
    ws = _lib.WORKSPACE.get("location", lib.kpms_resample_location_workspace_bytes(N, T, dk), dev)
    _lib.check(lib.kpms_resample_location(
        _lib.ptr(Y), _lib.ptr(mask), _lib.ptr(h), _lib.ptr(x), _lib.ptr(s), _lib.ptr(Cdfull),
        _lib.ptr(sigmasq), float(sigmasq_loc), _lib.ptr(w), _lib.ptr(w_last), N, T, k, dk, D,
        _lib.ptr(vout), _lib.ptr(loglik), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)),
        "kpms_resample_location")
    if return_info:
        return vout, {"loglik": loglik}
    return vout


# ------------------------------------------------------------------------ a13
def obs_variance_stats(*, Y, mask, Cd, v, h, x, s, Gamma=None, **kwargs):
    """gibbs.py:86-94 -> (S_y (k,) float64, mask.sum() (1,) float64)."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    mask, v, h, x, s = f32(mask, dev), f32(v, dev), f32(h, dev), f32(x, dev), f32(s, dev)
    D = x.shape[-1]
    Cdfull = expand_Cd(f64(Cd, dev), _gamma(Gamma, k, dev), dk)
    S_y = torch.empty((k,), dtype=torch.float64, device=dev)
    msum = torch.empty((1,), dtype=torch.float64, device=dev)
    ws = _lib.WORKSPACE.get("obsvar", lib.kpms_obs_variance_stats_workspace_bytes(N, T, k), dev)
    _lib.check(lib.kpms_obs_variance_stats(
        _lib.ptr(Y), _lib.ptr(mask), _lib.ptr(v), _lib.ptr(h), _lib.ptr(x), _lib.ptr(s), _lib.ptr(Cdfull),
        N, T, k, dk, D, _lib.ptr(S_y), _lib.ptr(msum), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)),
        "kpms_obs_variance_stats")
    return S_y, msum


# ------------------------------------------------------------------------ (f)2: diagnostics
def log_joint_likelihood(*, Y, mask, x, s, v, h, z, pi, Ab, Q, Cd, sigmasq, sigmasq_loc, nu_s, s_0,
                         Gamma=None, **kwargs):
    """util.py:266-275 -> dict {'Y','x','z','v','s'} of float64 scalars (device tensors)."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    mask, x, s, v, h = f32(mask, dev), f32(x, dev), f32(s, dev), f32(v, dev), f32(h, dev)
    z = i32(z, dev)
    pi, Ab, Q, sigmasq = f64(pi, dev), f64(Ab, dev), f64(Q, dev), f64(sigmasq, dev)
    K, D, L = _dims(Ab)
    Cdfull = expand_Cd(f64(Cd, dev), _gamma(Gamma, k, dev), dk)
    Qinv = spd_inv(Q).contiguous()                              # util.py:237-238
    logdet = torch.linalg.slogdet(Q)[1].contiguous()
    out = torch.empty((5,), dtype=torch.float64, device=dev)
    ws = _lib.WORKSPACE.get("logjoint", lib.kpms_log_joint_workspace_bytes(), dev)
    _lib.check(lib.kpms_log_joint(
        _lib.ptr(Y), _lib.ptr(mask), _lib.ptr(x), _lib.ptr(s), _lib.ptr(v), _lib.ptr(h), _lib.ptr(z), _lib.ptr(pi),
        _lib.ptr(Ab), _lib.ptr(Qinv), _lib.ptr(logdet), _lib.ptr(Cdfull), _lib.ptr(sigmasq), float(sigmasq_loc),
        float(nu_s), float(s_0), N, T, k, dk, D, L, K, _lib.ptr(out), _lib.ptr(ws), ws.numel(),
        _lib.stream_ptr(dev)), "kpms_log_joint")
    return {name: out[i] for i, name in enumerate(("Y", "x", "z", "v", "s"))}


def stateseq_stats(stateseqs, mask):
    """util.py:41-45 -> (usage (max state + 1,), durations of the runs of the flattened unmasked
    sequence), both int64 device tensors.  Index bookkeeping only (no kernel of its own)."""
    z = stateseqs
    m = mask[..., -z.shape[-1]:]
    sq = z.reshape(-1)[m.reshape(-1) > 0].to(torch.int64)
    usage = torch.bincount(sq)
    durations = torch.unique_consecutive(sq, return_counts=True)[1]
    return usage, durations


# ------------------------------------------------------------------------ (f)1
def obs_params_stats(*, Y, mask, sigmasq, v, h, x, s, **kwargs):
    """Per-keypoint weighted moments of gibbs.py:58-82 -> (W (k,D+1,D+1), V (k,D+1,dk)) float64."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    mask, v, h, x, s = f32(mask, dev), f32(v, dev), f32(h, dev), f32(x, dev), f32(s, dev)
    sigmasq = f64(sigmasq, dev)
    D = x.shape[-1]
    W = torch.empty((k, D + 1, D + 1), dtype=torch.float64, device=dev)
    V = torch.empty((k, D + 1, dk), dtype=torch.float64, device=dev)
    ws = _lib.WORKSPACE.get("obsparams", lib.kpms_obs_params_stats_workspace_bytes(N, T, k, dk, D), dev)
    _lib.check(lib.kpms_obs_params_stats(
        _lib.ptr(Y), _lib.ptr(mask), _lib.ptr(v), _lib.ptr(h), _lib.ptr(x), _lib.ptr(s), _lib.ptr(sigmasq),
        N, T, k, dk, D, _lib.ptr(W), _lib.ptr(V), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)),
        "kpms_obs_params_stats")
    return W, V


def sample_obs_params(key, W, V, Gamma, *, sigmasq_C, draws=None):
    """gibbs.py:69-82 given the (possibly all-reduced) moments -> Cd ((k-1) dk, D+1) float64."""
    dev = W.device
    k, D1, dk = V.shape
    Gamma = f64(Gamma, dev)
    S4 = torch.einsum('ji,jl,jab->aibl', Gamma, Gamma, W)                       # gibbs.py:69-74
    S_xx = torch.kron(S4.reshape(D1 * (k - 1), D1 * (k - 1)).contiguous(),
                      torch.eye(dk, dtype=torch.float64, device=dev))
    S_yx = torch.einsum('ji,jac->aic', Gamma, V).reshape(-1)                    # gibbs.py:77-80
    n = dk * D1 * (k - 1)
    Sigma_n = spd_inv(torch.eye(n, dtype=torch.float64, device=dev) / sigmasq_C + S_xx)
    Sigma_n = 0.5 * (Sigma_n + Sigma_n.T)
    mu_n = Sigma_n @ S_yx
    if draws is not None:
        w = f64(draws["w"], dev)
    else:
        w = torch.randn((n,), dtype=torch.float64, device=dev, generator=dist.torch_generator(dist.seed_of(key), 13, dev))
    draw = mu_n + torch.linalg.cholesky_ex(Sigma_n)[0] @ w                            # jr.multivariate_normal
    return draw.reshape(D1, dk * (k - 1)).T.contiguous()


def resample_obs_params(key, *, Y, mask, sigmasq, v, h, x, s, sigmasq_C, Gamma=None, draws=None, **kwargs):
    """gibbs.py:58-82 -> Cd ((k-1) dk, D+1) float64."""
    W, V = obs_params_stats(Y=Y, mask=mask, sigmasq=sigmasq, v=v, h=h, x=x, s=s)
    k = Y.shape[2]
    return sample_obs_params(key, W, V, _gamma(Gamma, k, W.device), sigmasq_C=sigmasq_C, draws=draws)


def sample_obs_variance(key, S_y, msum, *, nu_sigma, sigmasq_0, draws=None):
    """gibbs.py:94-96 given the (possibly all-reduced) statistics."""
    variance = (nu_sigma * sigmasq_0 + S_y) / (nu_sigma + 3 * msum)   # literal 3: gibbs.py:94
    degs = (nu_sigma + 3 * msum) * torch.ones_like(variance)
    if draws is not None:
        g = f64(draws["g"], S_y.device)
    else:
        g = dist.sample_gamma(dist.torch_generator(dist.seed_of(key), 13, S_y.device), degs / 2)
    return dist.sample_scaled_inv_chi2(degs, variance, g)


def resample_obs_variance(key, *, Y, mask, Cd, v, h, x, s, nu_sigma, sigmasq_0, Gamma=None, draws=None,
                          **kwargs):
    """gibbs.py:84-96 -> sigmasq (k,) float64."""
    S_y, msum = obs_variance_stats(Y=Y, mask=mask, Cd=Cd, v=v, h=h, x=x, s=s, Gamma=Gamma)
    return sample_obs_variance(key, S_y, msum, nu_sigma=nu_sigma, sigmasq_0=sigmasq_0, draws=draws)


# ------------------------------------------------------------------------ a14
def resample_scales(key, *, x, v, h, Y, Cd, sigmasq, nu_s, s_0, Gamma=None, draws=None, session_offset=0,
                    mask=None, return_stats=False, **kwargs):
    """gibbs.py:117-125 -> s (N,T,k) float32 (no mask, as the reference).
    Without `draws` the Gamma variates are generated inside the kernel (keyed by the global element
    index).  `return_stats=True` (needs `mask`) also returns (S_y (k,), mask.sum() (1,)) of
    gibbs.py:86-94 for the state just produced — what the next sweep's resample_obs_variance needs —
    accumulated in the same pass over Y."""
    lib = _lib.load()
    Y = f32(Y)
    dev = Y.device
    N, T, k, dk = Y.shape
    x, v, h, sigmasq = f32(x, dev), f32(v, dev), f32(h, dev), f64(sigmasq, dev)
    D = x.shape[-1]
    Cdfull = expand_Cd(f64(Cd, dev), _gamma(Gamma, k, dev), dk)
    g = f32(draws["g"], dev) if draws is not None else None
    sout = torch.empty((N, T, k), dtype=torch.float32, device=dev)
    S_y = msum = ws = None
    if return_stats:
        if mask is None:
            raise ValueError("resample_scales(return_stats=True) needs mask")
        mask = f32(mask, dev)
        S_y = torch.empty((k,), dtype=torch.float64, device=dev)
        msum = torch.empty((1,), dtype=torch.float64, device=dev)
        ws = _lib.WORKSPACE.get("obsvar", lib.kpms_obs_variance_stats_workspace_bytes(N, T, k), dev)
    _lib.check(lib.kpms_resample_scales_stats(
        _lib.ptr(Y), _lib.ptr(mask) if return_stats else None, _lib.ptr(v), _lib.ptr(h), _lib.ptr(x),
        _lib.ptr(Cdfull), _lib.ptr(sigmasq), float(nu_s), float(s_0), _lib.ptr(g), dist.seed_of(key),
        dist.STREAM_S, int(session_offset) * T * k, N, T, k, dk, D, _lib.ptr(sout), _lib.ptr(S_y),
        _lib.ptr(msum), _lib.ptr(ws), ws.numel() if ws is not None else 0, _lib.stream_ptr(dev)),
        "kpms_resample_scales_stats")
    if return_stats:
        return sout, (S_y, msum)
    return sout


# ------------------------------------------------------------------------ a8-a9
def ar_sufficient_stats(*, nlags, num_states, mask, x, z, **kwargs):
    """gibbs.py:129-131,141-143 -> F (K,P,P) float64 over f = [x_in, 1, x_out]."""
    lib = _lib.load()
    x = f32(x)
    dev = x.device
    mask, z = f32(mask, dev), i32(z, dev)
    N, T, D = x.shape
    L, K = int(nlags), int(num_states)
    P = D * L + 1 + D
    F = torch.empty((K, P, P), dtype=torch.float64, device=dev)
    ws = _lib.WORKSPACE.get("arstats", lib.kpms_ar_stats_workspace_bytes(N, T, D, L, K), dev)
    _lib.check(lib.kpms_ar_stats(_lib.ptr(x), _lib.ptr(z), _lib.ptr(mask), N, T, D, L, K, _lib.ptr(F),
                                 _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev)), "kpms_ar_stats")
    return F


def sample_ar_params(key, F, *, nlags, nu_0, S_0, M_0, K_0, draws=None):
    """gibbs.py:132-136 + distributions.py:23-42 from the (possibly all-reduced) statistics."""
    dev = F.device
    K, P, _ = F.shape
    D = S_0.shape[-1] if torch.is_tensor(S_0) else len(S_0)
    P1 = D * int(nlags) + 1
    S_0, M_0, K_0 = f64(S_0, dev), f64(M_0, dev), f64(K_0, dev)
    S_in_in, S_out_in, S_out_out = F[:, :P1, :P1], F[:, P1:, :P1], F[:, P1:, P1:]
    count = F[:, P1 - 1, P1 - 1]
    K_0_inv = spd_inv(K_0)
    K_n_inv = K_0_inv + S_in_in                     # gibbs.py:133 inverts this and :135 inverts it back
    K_n = spd_inv(K_n_inv)
    M_n = (M_0 @ K_0_inv + S_out_in) @ K_n
    S_n = S_0 + S_out_out + (M_0 @ K_0_inv @ M_0.T - M_n @ K_n_inv @ M_n.transpose(-1, -2))
    nu = nu_0 + count
    if draws is not None:
        chi2, normals, G = f64(draws["chi2"], dev), f64(draws["normals"], dev), f64(draws["G"], dev)
    else:
        gen = dist.torch_generator(dist.seed_of(key), 11, dev)
        dof = nu[:, None] - torch.arange(D, dtype=torch.float64, device=dev)[None, :]
        chi2 = dist.sample_gamma(gen, dof / 2) * 2                     # distributions.py:17-18
        normals = torch.randn((K, D * (D - 1) // 2), dtype=torch.float64, device=dev, generator=gen)
        G = torch.randn((K, D, P1), dtype=torch.float64, device=dev, generator=gen)
    Ab, Q = dist.sample_mniw(nu, S_n, M_n, K_n, chi2, normals, G)
    return Ab.contiguous(), Q.contiguous()


def resample_ar_params(key, *, nlags, num_states, mask, x, z, nu_0, S_0, M_0, K_0, draws=None, **kwargs):
    """gibbs.py:139-145 -> (Ab (K,D,D*L+1), Q (K,D,D)) float64."""
    F = ar_sufficient_stats(nlags=nlags, num_states=num_states, mask=mask, x=x, z=z)
    return sample_ar_params(key, F, nlags=nlags, nu_0=nu_0, S_0=S_0, M_0=M_0, K_0=K_0, draws=draws)


def resample_regression_params(key, mask, x_in, x_out, nu_0, S_0, M_0, K_0, draws=None):
    """gibbs.py:127-136 for one state with explicit regressors (small-problem utility; the
    sweep uses `resample_ar_params`, which never materialises x_in)."""
    x_in, x_out = f64(x_in), f64(x_out, x_in.device)
    m = f64(mask, x_in.device)
    f = torch.cat([x_in, x_out], dim=1) * torch.sqrt(m)[:, None]
    P1 = x_in.shape[1]
    F = (f.T @ f)[None]
    F[0, P1 - 1, P1 - 1] = m.sum()
    nl = (P1 - 1) // x_out.shape[1]
    Ab, Q = sample_ar_params(key, F, nlags=nl, nu_0=nu_0, S_0=S_0, M_0=M_0, K_0=K_0, draws=draws)
    return Ab[0], Q[0]


# ------------------------------------------------------------------------ a10-a12
def transition_counts(*, z, mask, num_states, **kwargs):
    """util.py:178-189 on mask[..., -z.shape[-1]:] (gibbs.py:149) -> (K,K) int64."""
    lib = _lib.load()
    z = i32(z)
    dev = z.device
    mask = f32(mask, dev)
    N, Tm = z.shape
    T = mask.shape[-1]
    counts = torch.empty((num_states, num_states), dtype=torch.int64, device=dev)
    mview = mask[:, T - Tm:]
    _lib.check(lib.kpms_count_transitions(_lib.ptr(z), _lib.c_void_p(mview.data_ptr()), T, N, Tm,
                                          int(num_states), _lib.ptr(counts), _lib.stream_ptr(dev)),
               "kpms_count_transitions")
    return counts


def sample_ms(key, counts, betas, alpha, kappa, draws=None, max_customers=None, thin=False):
    """transitions.py:23-33 -> (ms (K,K) int64, ms_raw (K,K) int64 = the CRP table counts).
    thin=False (default): ms = ms_raw, which is what the reference returns — its binomial thinning of
    the sticky diagonal (transitions.py:28-31) is assigned to a temporary copy (`newms.flat[::K+1]`
    slices a flat iterator) and never reaches the result; pinned by
    tests/golden/reference_sweep_bench.npz.  thin=True: the thinning the code intends, as the
    Bernoulli sum over uniforms.  Without `draws` the uniforms are generated inside the kernel (Philox,
    customer c of cell (i,j) = number starts[i,j]+c of the stream): nothing of the size of the data
    is materialised and every rank of a sharded job gets the same counts.  With injected `draws`
    (`randseq`, `u_thin`) a capacity overflow raises.  `max_customers` is accepted for
    compatibility and ignored."""
    lib = _lib.load()
    dev = counts.device
    K = counts.shape[0]
    counts = counts.to(torch.int64).contiguous()
    betas = f64(betas, dev)
    flat = counts.flatten()
    starts = (torch.cumsum(flat, 0) - flat).reshape(K, K).contiguous()   # transitions.py:12-15
    ms = torch.empty((K, K), dtype=torch.int64, device=dev)
    ms_raw = torch.empty((K, K), dtype=torch.int64, device=dev)
    if draws is None or "randseq" not in draws:
        seed = dist.seed_of(key)
        _lib.check(lib.kpms_crp_tablecounts_philox(
            _lib.ptr(counts), _lib.ptr(starts), _lib.ptr(betas), float(alpha), float(kappa), seed,
            dist.STREAM_CRP, dist.STREAM_THIN, K, _lib.ptr(ms), _lib.ptr(ms_raw), _lib.stream_ptr(dev)),
            "kpms_crp_tablecounts_philox")
        return (ms if thin else ms_raw), ms_raw
    randseq = f32(draws["randseq"], dev)
    u_thin = f32(draws["u_thin"], dev) if thin else torch.zeros((K, 1), dtype=torch.float32, device=dev)
    nrand = randseq.numel()
    if nrand == 0:
        randseq = torch.zeros(1, dtype=torch.float32, device=dev)
    overflow = torch.zeros((1,), dtype=torch.int32, device=dev)
    _lib.check(lib.kpms_crp_tablecounts(
        _lib.ptr(counts), _lib.ptr(starts), _lib.ptr(betas), float(alpha), float(kappa), _lib.ptr(randseq),
        nrand, _lib.ptr(u_thin), int(u_thin.shape[1]), K, _lib.ptr(ms), _lib.ptr(ms_raw), _lib.ptr(overflow),
        _lib.stream_ptr(dev)), "kpms_crp_tablecounts")
    flag = int(overflow.item()) & (3 if thin else 2)
    if flag:
        raise _lib.KpmsError("sample_ms: injected draws too short (%s)" % ", ".join(
            n for b, n in ((1, "u_thin has fewer columns than tables to thin"),
                           (2, "randseq shorter than the number of transitions")) if flag & b))
    return (ms if thin else ms_raw), ms_raw


def sample_hdp_transitions(key, counts, betas, alpha, kappa, gamma, draws=None, max_customers=None, thin=False,
                           return_info=False):
    """transitions.py:35-40 -> (betas (K,), pi (K,K)) float64.  With `return_info=True` also a dict
    with the table counts `ms` and the two Dirichlet concentrations."""
    dev = counts.device
    K = counts.shape[0]
    ms, _ = sample_ms(key, counts, betas, alpha, kappa, draws=draws, thin=thin)
    gen = dist.torch_generator(dist.seed_of(key), 12, dev)
    if draws is not None and "betas_g" in draws:
        gb = f64(draws["betas_g"], dev)
        betas_new = gb / gb.sum()
    else:
        betas_new = dist.sample_dirichlet(gen, ms.sum(0).to(torch.float64) + gamma / K)
    conc = alpha * betas_new[None, :] + counts.to(torch.float64) + kappa * torch.eye(
        K, dtype=torch.float64, device=dev)
    if draws is not None and "pi_g" in draws:
        gp = f64(draws["pi_g"], dev)
        pi = gp / gp.sum(-1, keepdim=True)
    else:
        pi = dist.sample_dirichlet(gen, conc)
    if return_info:
        return betas_new, pi, {"ms": ms, "betas_conc": ms.sum(0).to(torch.float64) + gamma / K, "pi_conc": conc}
    return betas_new, pi


def sample_transitions(key, counts, alpha, kappa, draws=None):
    """transitions.py:42-44 -> pi (K,K) float64; `draws={'pi_g': Gamma(conc) variates}` injects the draw."""
    dev = counts.device
    K = counts.shape[0]
    conc = counts.to(torch.float64) + alpha + kappa * torch.eye(K, dtype=torch.float64, device=dev)
    if draws is not None and "pi_g" in draws:
        gp = f64(draws["pi_g"], dev)
        return gp / gp.sum(-1, keepdim=True)
    return dist.sample_dirichlet(dist.torch_generator(dist.seed_of(key), 14, dev), conc)


def resample_hdp_transitions(key, *, z, mask, betas, alpha, kappa, gamma, num_states, draws=None, thin=False,
                             return_info=False, **kwargs):
    """gibbs.py:148-151 -> (betas, pi).  No device->host copy of z / mask: the histogram and
    the CRP table counts run on the GPU (the reference round-trips through numba)."""
    counts = transition_counts(z=z, mask=mask, num_states=num_states)
    return sample_hdp_transitions(key, counts, betas, alpha, kappa, gamma, draws=draws, thin=thin,
                                  return_info=return_info)


def resample_transitions(key, *, z, mask, alpha, kappa, num_states, draws=None, **kwargs):
    """gibbs.py:153-156 -> pi."""
    counts = transition_counts(z=z, mask=mask, num_states=num_states)
    return sample_transitions(key, counts, alpha, kappa, draws=draws)
