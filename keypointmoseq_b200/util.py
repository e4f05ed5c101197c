"""Helpers of the reference's `keypoint_moseq/util.py` on torch tensors (same names, argument
meaning and return values).

The kernels never materialise lag windows, rotation matrices or `kron(Gamma, I)`; these functions
exist for API parity, for the initialisers and for small host-side algebra.  They work on whatever
device their inputs live on.  `count_transitions` calls the CUDA histogram kernel;
`merge_data` / `interpolate_keypoints` are the ingest layer (host arrays in, device tensors out).
Each function is pinned against the reference's own implementation in
tests/golden/reference_util.npz (tests/test_util_initialize.py).
"""
import functools
import math

import numpy as np
import torch

from . import _lib


def _cast(t, dtype, device):
    t = torch.as_tensor(t)
    if device is not None and t.device != torch.device(device):
        t = t.to(device)
    return t.to(dtype).contiguous()


def f32(t, device=None):
    return _cast(t, torch.float32, device)


def f64(t, device=None):
    return _cast(t, torch.float64, device)


def i32(t, device=None):
    return _cast(t, torch.int32, device)


@functools.lru_cache(maxsize=None)
def _center_embedding_np(k):
    return np.linalg.svd(np.eye(k) - np.ones((k, k)) / k)[0][:, :-1]


def center_embedding(k, device="cuda"):
    """util.py:38-39.  The basis of the sum-zero subspace is arbitrary; it is computed once
    per k on the host (LAPACK SVD, as the reference) and cached, so every step of a fit
    sees the same Gamma.  Pass `Gamma=` explicitly to the gibbs steps to override."""
    return torch.as_tensor(_center_embedding_np(int(k)), dtype=torch.float64, device=device)


def expand_Cd(Cd, Gamma, dk):
    """gibbs.py:14: kron(Gamma, I_dk) @ Cd -> (k*dk, D+1) float64."""
    Cd = f64(Cd)
    Gamma = f64(Gamma, Cd.device)
    k, km1 = Gamma.shape
    # (kron(Gamma, I) @ Cd)[(j, a), :] = sum_i Gamma[j, i] Cd[(i, a), :]
    return torch.einsum("ji,iac->jac", Gamma, Cd.reshape(km1, dk, -1)).reshape(k * dk, -1).contiguous()


def pad_affine(x):
    """util.py:169-176: append a constant-one coordinate."""
    return torch.nn.functional.pad(x, (0, 1), value=1.0)


def get_lags(x, nlags):
    """util.py:192-211: row i of the result is the window [x[i], ..., x[i+nlags-1]] (oldest first),
    i.e. the regressors of frame i + nlags; shape (..., T - nlags, nlags * D)."""
    win = x.unfold(-2, nlags, 1)[..., :-1, :, :]          # (..., T-nlags, D, nlags)
    return win.transpose(-1, -2).reshape(*win.shape[:-2], -1)


def ar_to_lds(As, bs, Qs, Cs):
    """util.py:214-230: companion-form LDS of an order-l vector autoregression.
    As (K, D, D*l), bs (K, D), Qs (K, D, D), Cs (p, D)  ->  A_ (K, n, n), b_ (K, n), Q_ (K, n, n),
    C_ (p, n), n = D*l: the state stacks the l most recent x (oldest first), the top n-D rows shift
    it, the last D rows apply the regression; lag copies carry 1e-2 process noise."""
    K, D, n = As.shape
    o = dict(dtype=As.dtype, device=As.device)
    shift = torch.eye(n, **o)[D:]                                         # (n-D, n): picks blocks 1..l-1
    A_ = torch.cat([shift.expand(K, n - D, n), As], dim=1)
    b_ = torch.cat([torch.zeros((bs.shape[0], n - D), **o), bs], dim=1)
    Q_ = torch.zeros((Qs.shape[0], n, n), **o)
    Q_.diagonal(dim1=-2, dim2=-1)[:, :n - D] = 1e-2
    Q_[:, n - D:, n - D:] = Qs
    C_ = torch.cat([torch.zeros((Cs.shape[0], n - D), **o), Cs], dim=1)
    return A_, b_, Q_, C_


def vector_to_angle(V):
    """util.py:66-82: angle of a planar vector in [-pi, pi), with the reference's 1e-10 offset on
    the first coordinate (the cut is at -pi: (-1, 0) maps to -pi)."""
    a = torch.atan2(V[..., 1], V[..., 0] + 1e-10)
    return torch.remainder(a + math.pi, 2 * math.pi) - math.pi


def angle_to_rotation_matrix(h, keypoint_dim=3):
    """util.py:85-109: rotation by h about the third axis (identity beyond the first two)."""
    c, s = torch.cos(h), torch.sin(h)
    m = torch.zeros((*h.shape, keypoint_dim, keypoint_dim), dtype=h.dtype, device=h.device)
    m.diagonal(dim1=-2, dim2=-1)[...] = 1.0
    m[..., 0, 0], m[..., 0, 1], m[..., 1, 0], m[..., 1, 1] = c, -s, s, c
    return m


def _rotate_plane(Y, c, s):
    out = Y.clone()
    out[..., 0] = c * Y[..., 0] - s * Y[..., 1]
    out[..., 1] = s * Y[..., 0] + c * Y[..., 1]
    return out


def affine_transform(Y, v, h):
    """util.py:112-138: R(h) Y + v for Y (..., k, d), v (..., d), h (...)."""
    return _rotate_plane(Y, torch.cos(h)[..., None], torch.sin(h)[..., None]) + v[..., None, :]


def inverse_affine_transform(Y, v, h):
    """util.py:140-166: R(-h) (Y - v)."""
    return _rotate_plane(Y - v[..., None, :], torch.cos(h)[..., None], -torch.sin(h)[..., None])


def expected_keypoints(*, Y, v, h, x, Cd, Gamma=None, **kwargs):
    """util.py:12-17: R(h) Gamma (Cd [x; 1]) + v."""
    k, d = Y.shape[-2:]
    Gamma = center_embedding(k, Y.device).to(x.dtype) if Gamma is None else torch.as_tensor(Gamma, device=Y.device).to(x.dtype)
    low = (pad_affine(x) @ torch.as_tensor(Cd, device=x.device).to(x.dtype).T).reshape(*Y.shape[:-2], k - 1, d)
    return affine_transform(torch.einsum("ji,...id->...jd", Gamma, low), v, h)


def pca_transform(components, means, *, Y, v, h, **kwargs):
    """util.py:20-22."""
    y = inverse_affine_transform(Y, v, h).reshape(*Y.shape[:-2], -1)
    return (y - means) @ components


def ensure_symmetric(X):
    """util.py:62-64."""
    return 0.5 * (X + X.transpose(-1, -2))


def stateseq_stats(stateseqs, mask):
    """util.py:41-45 -> (usage, durations) of the flattened unmasked state sequence."""
    from .gibbs import stateseq_stats as _impl
    return _impl(stateseqs, mask)


def count_transitions(num_states, mask, stateseqs):
    """util.py:178-189 -> (K,K) float64 counts (CUDA integer histogram, bit-exact)."""
    lib = _lib.load()
    z = i32(stateseqs)
    dev = z.device
    m = f32(mask, dev)
    N, Tm = z.shape
    if m.shape != (N, Tm):
        raise ValueError("mask must have the shape of stateseqs (pass mask[..., -z.shape[-1]:])")
    counts = torch.empty((num_states, num_states), dtype=torch.int64, device=dev)
    _lib.check(lib.kpms_count_transitions(_lib.ptr(z), _lib.ptr(m), Tm, N, Tm, int(num_states),
                                          _lib.ptr(counts), _lib.stream_ptr(dev)),
               "kpms_count_transitions")
    return counts.to(torch.float64)


# ------------------------------------------------------------------------------- ingest
def merge_data(data_dict, keys=None, batch_length=None, device=None, dtype=torch.float32):
    """util.py:47-60: stack variable-length sessions {name: array (T_i, ...)} into one zero-padded
    batch.  Sessions longer than `batch_length` (default: the longest session) are cut into
    consecutive rows of that length.  Returns (data (rows, batch_length, ...), mask (rows,
    batch_length) with 1 on real frames, keys [(name, piece index), ...]).

    The batch is assembled directly in its final (pinned, when `device` is a CUDA device) buffer,
    one copy per session, and uploaded once; with device=None NumPy arrays are returned like the
    reference."""
    names = sorted(data_dict.keys()) if keys is None else list(keys)
    lengths = [int(np.shape(data_dict[n])[0]) for n in names]
    if batch_length is None:
        batch_length = max(lengths)
    batch_length = int(batch_length)
    pieces = [-(-t // batch_length) for t in lengths]
    rows = int(sum(pieces))
    trailing = tuple(np.shape(data_dict[names[0]])[1:])
    np_dtype = np.float64 if device is None else torch.empty((), dtype=dtype).numpy().dtype
    to_cuda = device is not None and torch.device(device).type == "cuda"
    if to_cuda:
        data_t = torch.zeros((rows * batch_length, *trailing), dtype=dtype).pin_memory()
        mask_t = torch.zeros((rows * batch_length,), dtype=dtype).pin_memory()
        data, mask = data_t.numpy(), mask_t.numpy()
    else:
        data = np.zeros((rows * batch_length, *trailing), dtype=np_dtype)
        mask = np.zeros((rows * batch_length,), dtype=np_dtype)
    row0, out_keys = 0, []
    for name, t, p in zip(names, lengths, pieces):
        start = row0 * batch_length
        data[start:start + t] = np.asarray(data_dict[name])
        mask[start:start + t] = 1
        out_keys.extend((name, i) for i in range(p))
        row0 += p
    data = data.reshape(rows, batch_length, *trailing)
    mask = mask.reshape(rows, batch_length)
    if device is None:
        return data, mask, out_keys
    if to_cuda:
        return (data_t.reshape(data.shape).to(device, non_blocking=True),
                mask_t.reshape(mask.shape).to(device, non_blocking=True), out_keys)
    return torch.as_tensor(data, device=device), torch.as_tensor(mask, device=device), out_keys


def interpolate_keypoints(keypoints, outliers, axis=1):
    """util.py:23-35: replace flagged detections by linear interpolation along time between the
    nearest unflagged frames (constant extrapolation at the ends), independently per keypoint.
    `keypoints` (..., T at `axis`, ..., d), `outliers` the same shape without the last axis.
    Vectorised (searchsorted over the good frames of every track at once); torch in, torch out —
    NumPy in, NumPy out."""
    as_numpy = not torch.is_tensor(keypoints)
    kp = torch.as_tensor(keypoints)
    bad = torch.as_tensor(outliers, device=kp.device).to(torch.bool)
    kp = kp.movedim(axis, 0)
    bad = bad.movedim(axis, 0)
    T = kp.shape[0]
    shape = kp.shape
    d = shape[-1]
    val = kp.reshape(T, -1).to(torch.float64)                                # (T, tracks*d)
    bad = bad.reshape(T, -1).repeat_interleave(d, dim=1)                     # the flag covers all d coordinates
    tt = torch.arange(T, device=kp.device)[:, None].expand_as(bad)
    big = T + 1
    # index of the nearest good frame at or before t (-1: none) and at or after t (T: none)
    prev = torch.where(bad, torch.full_like(tt, -1), tt).cummax(0).values
    nxt = torch.where(bad, torch.full_like(tt, big), tt).flip(0).cummin(0).values.flip(0)
    has_prev, has_next = prev >= 0, nxt < big
    lo = torch.where(has_prev, prev, nxt).clamp(0, T - 1)
    hi = torch.where(has_next, nxt, prev).clamp(0, T - 1)
    vlo, vhi = val.gather(0, lo), val.gather(0, hi)
    span = (hi - lo).clamp(min=1).to(torch.float64)
    wgt = ((tt - lo).to(torch.float64) / span).clamp(0.0, 1.0)
    out = vlo + (vhi - vlo) * wgt
    out = torch.where(bad, out, val).to(kp.dtype).reshape(shape).movedim(0, axis)
    return out.numpy() if as_numpy else out
