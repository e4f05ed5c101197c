"""Layout helpers of the reference's `keypoint_moseq/util.py`, on torch CUDA tensors.

These are the cheap index/shape utilities around the hot path (the kernels never
materialise `get_lags` or `kron(Gamma, I)`; they are provided for API parity and for
the host-side small math).  `count_transitions` calls the CUDA histogram kernel.
"""
import functools

import numpy as np
import torch

from . import _lib


def f32(t, device=None):
    t = torch.as_tensor(t)
    if device is not None and t.device != torch.device(device):
        t = t.to(device)
    return t.to(torch.float32).contiguous()


def f64(t, device=None):
    t = torch.as_tensor(t)
    if device is not None and t.device != torch.device(device):
        t = t.to(device)
    return t.to(torch.float64).contiguous()


def i32(t, device=None):
    t = torch.as_tensor(t)
    if device is not None and t.device != torch.device(device):
        t = t.to(device)
    return t.to(torch.int32).contiguous()


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
    eye = torch.eye(dk, dtype=torch.float64, device=Cd.device)
    return (torch.kron(Gamma, eye) @ Cd).contiguous()


def pad_affine(x):
    """util.py:169-176."""
    return torch.cat([x, torch.ones((*x.shape[:-1], 1), dtype=x.dtype, device=x.device)], dim=-1)


def get_lags(x, nlags):
    """util.py:192-211: row t = [x[t-nlags], ..., x[t-1]] for t = nlags..T-1."""
    T = x.shape[-2]
    return torch.cat([x[..., i:T - nlags + i, :] for i in range(nlags)], dim=-1)


def ar_to_lds(As, bs, Qs, Cs):
    """util.py:214-230."""
    k, l = As.shape[1], As.shape[2] // As.shape[1]
    kw = dict(dtype=As.dtype, device=As.device)
    A_ = torch.zeros((As.shape[0], k * l, k * l), **kw)
    A_[:, :-k, k:] = torch.eye(k * (l - 1), **kw)
    A_[:, -k:] = As
    Q_ = torch.zeros((Qs.shape[0], k * l, k * l), **kw)
    Q_[:, :-k, :-k] = torch.eye(k * (l - 1), **kw) * 1e-2
    Q_[:, -k:, -k:] = Qs
    b_ = torch.zeros((bs.shape[0], k * l), **kw)
    b_[:, -k:] = bs
    C_ = torch.zeros((Cs.shape[0], k * l), **kw)
    C_[:, -k:] = Cs
    return A_, b_, Q_, C_


def vector_to_angle(V):
    """util.py:66-82."""
    y, x = V[..., 1], V[..., 0] + 1e-10
    return torch.remainder(torch.atan(y / x) + (x > 0) * np.pi, 2 * np.pi) - np.pi


def angle_to_rotation_matrix(h, keypoint_dim=3):
    """util.py:85-109."""
    m = torch.eye(keypoint_dim, dtype=h.dtype, device=h.device).repeat(*h.shape, 1, 1)
    m[..., 0, 0] = torch.cos(h)
    m[..., 1, 1] = torch.cos(h)
    m[..., 0, 1] = -torch.sin(h)
    m[..., 1, 0] = torch.sin(h)
    return m


def affine_transform(Y, v, h):
    """util.py:112-138: R(h) Y + v."""
    rot = angle_to_rotation_matrix(h, keypoint_dim=Y.shape[-1])
    return (Y[..., None, :] * rot[..., None, :, :]).sum(-1) + v[..., None, :]


def inverse_affine_transform(Y, v, h):
    """util.py:140-166: R(-h)(Y - v)."""
    rot = angle_to_rotation_matrix(-h, keypoint_dim=Y.shape[-1])
    return ((Y - v[..., None, :])[..., None, :] * rot[..., None, :, :]).sum(-1)


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


def merge_data(data_dict, keys=None, batch_length=None):
    """util.py:47-60 (host-side NumPy, as the reference)."""
    if keys is None:
        keys = sorted(data_dict.keys())
    max_length = np.max([data_dict[k].shape[0] for k in keys])
    if batch_length is None:
        batch_length = max_length

    def reshape(x):
        padding = (-x.shape[0]) % batch_length
        x = np.concatenate([x, np.zeros((padding, *x.shape[1:]))], axis=0)
        return x.reshape(-1, batch_length, *x.shape[1:])

    data = np.concatenate([reshape(data_dict[k]) for k in keys], axis=0)
    mask = np.concatenate([reshape(np.ones(data_dict[k].shape[0])) for k in keys], axis=0)
    keys = [(k, i) for k in keys for i in range(int(np.ceil(len(data_dict[k]) / batch_length)))]
    return data, mask, keys
