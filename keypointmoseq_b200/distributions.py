"""Samplers of the reference's `keypoint_moseq/distributions.py`.

Per-frame draws (uniforms for the state sequence, normals for the latents and the
location, gammas for the scales) come from the library's Philox kernels
(`kpms_fill_*`).  The handful of per-state conjugate draws (K inverse-Wishart /
matrix-normal pairs, K+1 Dirichlet vectors, k scaled-inv-chi2 variances) are tiny dense
float64 problems done with batched torch.linalg on the GPU.  Every sampler accepts
pre-generated variates so tests can inject the oracle's draws.
"""
import hashlib

import torch

from . import _lib

# stream ids keep the Philox streams of different draws of one sweep disjoint
STREAM_Z, STREAM_X, STREAM_X_LAST, STREAM_H, STREAM_V, STREAM_V_LAST, STREAM_S, STREAM_CRP, \
    STREAM_THIN = range(1, 10)


def seed_of(key):
    """Map the reference's `key` argument (an int, a tuple of ints, a torch.Generator, or
    anything hashable by repr) to a 64-bit Philox key."""
    if isinstance(key, torch.Generator):
        return int(torch.randint(0, 2 ** 62, (1,), generator=key).item())
    if isinstance(key, int):
        return key & (2 ** 64 - 1)
    if torch.is_tensor(key):
        key = tuple(int(v) for v in key.flatten().tolist())
    h = hashlib.blake2b(repr(key).encode(), digest_size=8).digest()
    return int.from_bytes(h, "little")


def torch_generator(seed, salt, device):
    g = torch.Generator(device=device)
    g.manual_seed((seed * 6364136223846793005 + salt) & (2 ** 63 - 1))
    return g


def _row_offset(shape, first_row):
    """Global element index of this shard's first element: the draws of row (session) n depend on
    its GLOBAL index first_row + n only, so results do not depend on how sessions are sharded
    (the reference splits one key per session: gibbs.py:20,54,112,145)."""
    per_row = 1
    for d in shape[1:]:
        per_row *= int(d)
    return int(first_row) * per_row


def fill_uniform(shape, seed, stream_id, device, first_row=0):
    out = torch.empty(shape, dtype=torch.float32, device=device)
    _lib.check(_lib.load().kpms_fill_uniform_at(_lib.ptr(out), out.numel(), _row_offset(shape, first_row), seed,
                                                stream_id, _lib.stream_ptr(device)), "kpms_fill_uniform_at")
    return out


def fill_normal(shape, seed, stream_id, device, first_row=0):
    out = torch.empty(shape, dtype=torch.float32, device=device)
    _lib.check(_lib.load().kpms_fill_normal_at(_lib.ptr(out), out.numel(), _row_offset(shape, first_row), seed,
                                               stream_id, _lib.stream_ptr(device)), "kpms_fill_normal_at")
    return out


def fill_gamma(shape, conc, seed, stream_id, device, first_row=0):
    out = torch.empty(shape, dtype=torch.float32, device=device)
    _lib.check(_lib.load().kpms_fill_gamma_at(_lib.ptr(out), out.numel(), _row_offset(shape, first_row),
                                              float(conc), seed, stream_id, _lib.stream_ptr(device)),
               "kpms_fill_gamma_at")
    return out


def sample_gamma(gen, a, b=1.0):
    """distributions.py:8-9 (float64 tensors)."""
    a = torch.as_tensor(a, dtype=torch.float64)
    return torch._standard_gamma(a, generator=gen) / b


def sample_scaled_inv_chi2(degs, variance, g):
    """distributions.py:14-15 with the Gamma(degs/2) variate supplied."""
    return 1.0 / (g / (degs * variance / 2))


def sample_dirichlet(gen, conc):
    """jr.dirichlet: normalised Gamma(conc, 1)."""
    g = torch._standard_gamma(conc.to(torch.float64), generator=gen)
    return g / g.sum(-1, keepdim=True)


def sample_invwishart(S, nu, chi2, normals):
    """distributions.py:29-37, batched over the leading dim.  `chi2` (...,n) ~
    chi2(nu - arange(n)); `normals` (...,n(n-1)/2) fill the strict upper triangle in
    row-major order."""
    n = S.shape[-1]
    chol = torch.linalg.cholesky_ex(S)[0]
    x = torch.diag_embed(torch.sqrt(chi2))
    iu = torch.triu_indices(n, n, 1, device=S.device)
    x[..., iu[0], iu[1]] = normals
    # qr(x,'r') of an upper-triangular x is x up to row signs, which cancel in T T^T
    T = torch.linalg.solve_triangular(x.transpose(-1, -2), chol.transpose(-1, -2), upper=False)
    T = T.transpose(-1, -2)
    return T @ T.transpose(-1, -2)


def sample_mn(M, U, V, G):
    """distributions.py:23-27."""
    return M + torch.linalg.cholesky_ex(U)[0] @ G @ torch.linalg.cholesky_ex(V)[0].transpose(-1, -2)


def sample_mniw(nu, S, M, K, chi2, normals, G):
    """distributions.py:39-42."""
    sigma = sample_invwishart(S, nu, chi2, normals)
    return sample_mn(M, sigma, K, G), sigma
