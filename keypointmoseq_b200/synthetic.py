"""Synthetic keypoint-SLDS data drawn from the model itself.

The reference's example pickles are stripped (`.MISSING_LARGE_BLOBS`), so every
config in BASELINE.json is realised as a seeded draw from the generative model
with the hyper-parameters of `examples/keypoint_slds.ipynb` cells 4 and 11.
NumPy only (host side); the arrays are moved to the GPU by the caller.
"""
import numpy as np


def center_embedding(k):
    """Orthonormal basis Gamma (k, k-1) of the sum-zero subspace (reference
    `util.py:38-39`).  The basis is arbitrary, so it is computed once here and
    passed explicitly to every consumer."""
    return np.linalg.svd(np.eye(k) - np.ones((k, k)) / k)[0][:, :-1]


def default_hypparams(D, L, K, *, full_model=True):
    """Hyper-parameter dicts with the reference's literal values
    (`keypoint_slds.ipynb` cells 4, 11; `depth_ar_hmm.ipynb` cell 4)."""
    if full_model:
        trans = dict(gamma=1e3, alpha=100.0, kappa=1e6 / 50, num_states=K)
        ar = dict(nu_0=D + 2, S_0=0.01 * np.eye(D),
                  M_0=np.pad(np.eye(D), ((0, 0), ((L - 1) * D, 1))),
                  K_0=10 * np.eye(D * L + 1), num_states=K, nlags=L)
    else:
        trans = dict(gamma=1e3, alpha=5.7, kappa=2e5, num_states=K)
        ar = dict(nu_0=D + 200, S_0=10 * np.eye(D),
                  M_0=np.pad(np.eye(D), ((0, 0), ((L - 1) * D, 1))),
                  K_0=0.1 * np.eye(D * L + 1), num_states=K, nlags=L)
    obs = dict(sigmasq_0=10.0, sigmasq_C=0.1, nu_sigma=1e5, nu_s=5.0, s_0=1.0)
    translation = dict(sigmasq_loc=0.5)
    return trans, ar, obs, translation


def make_params(rng, D, L, K, k, dk):
    """Model parameters: sticky transition matrix (self-probability ~0.95),
    near-identity AR dynamics on the newest lag, small process noise, random
    loading matrix."""
    pi = rng.dirichlet(np.full(K, 100.0 / K) + 1e-3, size=K) * 0.05 + 0.95 * np.eye(K)
    pi /= pi.sum(1, keepdims=True)
    Ab = np.zeros((K, D, D * L + 1))
    Ab[:, :, (L - 1) * D:L * D] = 0.9 * np.eye(D) + 0.05 * rng.standard_normal((K, D, D)) / np.sqrt(D)
    Ab[:, :, -1] = 0.1 * rng.standard_normal((K, D))
    W = rng.standard_normal((K, D, 2 * D)) * 0.05
    Q = 0.05 * np.eye(D) + W @ W.transpose(0, 2, 1) / (2 * D)
    Cd = np.concatenate([rng.standard_normal(((k - 1) * dk, D)) / np.sqrt(D),
                         rng.standard_normal(((k - 1) * dk, 1))], axis=1)
    betas = np.full(K, 1.0 / K)
    sigmasq = np.ones(k)
    return dict(pi=pi, Ab=Ab, Q=Q, Cd=Cd, betas=betas, sigmasq=sigmasq)


def simulate(seed, N, T, *, k=8, dk=2, D=10, L=3, K=50, ragged=False, dtype=np.float32, params=None):
    """Draw (data, states, params, Gamma) of the named shape.  `params`: model parameters to draw the
    sessions from (default: drawn from `seed` too) — shards of one sharded job share them.

    data   : Y (N,T,k,dk), mask (N,T) (trailing padding only when `ragged`)
    states : x (N,T,D), z (N,T-L), v (N,T,dk), h (N,T), s (N,T,k)
    params : Ab, Q, pi, betas, Cd, sigmasq (float64)
    """
    rng = np.random.default_rng(seed)
    p = make_params(rng, D, L, K, k, dk)     # always drawn, so that the session draws below do not depend on `params`
    if params is not None:
        p = {a: np.asarray(b) for a, b in params.items()}
    Gamma = center_embedding(k)
    # state sequences (frames L..T-1), vectorised over sessions
    z = np.empty((N, T - L), dtype=np.int64)
    cdf = np.cumsum(p['pi'], axis=1)
    cdf[:, -1] = 1.0
    cur = rng.integers(0, K, size=N)
    U = rng.random((N, T - L))
    for t in range(T - L):
        cur = (U[:, t, None] >= cdf[cur]).sum(1)
        cur = np.minimum(cur, K - 1)
        z[:, t] = cur
    # latents
    x = np.empty((N, T, D))
    x[:, :L] = rng.standard_normal((N, L, D))
    Lq = np.linalg.cholesky(p['Q'])
    E = rng.standard_normal((N, T - L, D))
    for t in range(L, T):
        zt = z[:, t - L]
        xin = np.concatenate([x[:, t - L:t].reshape(N, L * D), np.ones((N, 1))], axis=1)
        x[:, t] = np.einsum('nij,nj->ni', p['Ab'][zt], xin) + np.einsum('nij,nj->ni', Lq[zt], E[:, t - L])
    # pose
    h = np.cumsum(0.05 * rng.standard_normal((N, T)), axis=1)
    h = (h + np.pi) % (2 * np.pi) - np.pi
    v = np.cumsum(np.sqrt(0.5) * rng.standard_normal((N, T, dk)), axis=1)
    s = np.where(rng.random((N, T, k)) < 0.02, 10.0, 1.0)
    Ybar = Gamma @ (np.concatenate([x, np.ones((N, T, 1))], -1) @ p['Cd'].T).reshape(N, T, k - 1, dk)
    c, sn = np.cos(h), np.sin(h)
    Yrot = Ybar.copy()
    Yrot[..., 0] = c[..., None] * Ybar[..., 0] - sn[..., None] * Ybar[..., 1]
    Yrot[..., 1] = sn[..., None] * Ybar[..., 0] + c[..., None] * Ybar[..., 1]
    Y = Yrot + v[:, :, None, :] + rng.standard_normal((N, T, k, dk)) * np.sqrt(s * p['sigmasq'])[..., None]
    mask = np.ones((N, T))
    if ragged:
        lens = rng.integers(T // 2, T + 1, size=N)
        lens[0] = T
        mask = (np.arange(T)[None, :] < lens[:, None]).astype(np.float64)
        Y = Y * mask[:, :, None, None]   # merge_data pads with zeros (util.py:53-55)
    data = dict(Y=Y.astype(dtype), mask=mask.astype(dtype))
    states = dict(x=x.astype(dtype), z=z, v=v.astype(dtype), h=h.astype(dtype), s=s.astype(dtype))
    return data, states, p, Gamma


CONFIGS = {
    # BASELINE.json configs[0..4] -> synthetic stand-ins
    'cfg1': dict(N=1, T=36000, k=8, dk=2, D=7, L=3, K=100),
    'cfg2': dict(N=1, T=36000, k=8, dk=2, D=10, L=3, K=100),      # AR-HMM only
    'cfg3': dict(N=100, T=36000, k=8, dk=2, D=10, L=3, K=50),
    'cfg4': dict(N=1000, T=36000, k=16, dk=2, D=10, L=3, K=100),
    'cfg5': dict(N=32, T=31250, k=8, dk=2, D=10, L=3, K=100),     # K swept 20/100/500
}
