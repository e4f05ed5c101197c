"""Initialisers of the reference's `keypoint_moseq/initialize.py` (same names, keyword arguments and
return values) on torch tensors.

`initial_location`, `initial_heading` and `initial_latents` are frame-parallel torch algebra that
runs on the device the data lives on (one-off set-up, not part of the sweep); the PCA is fitted on
a random sample of frames as in the reference (initialize.py:24-25) but from the sample's
covariance matrix, on the device, instead of a host-side sklearn fit.  The parameter initialisers
draw from the priors through the same samplers the Gibbs steps use.
"""
import torch

from . import distributions as dist
from . import gibbs as G
from .util import center_embedding, f64, inverse_affine_transform, vector_to_angle


class PCAModel:
    """The part of sklearn.decomposition.PCA the reference uses (initialize.py:25-27 and the
    explained-variance plot of keypoint_slds.ipynb cell 7): `components_` (n_components, features),
    `mean_`, `explained_variance_`, `explained_variance_ratio_`, `transform`."""

    def __init__(self, n_components):
        self.n_components = int(n_components)

    def fit(self, X):
        X = X.to(torch.float64)
        n = X.shape[0]
        self.mean_ = X.mean(0)
        Xc = X - self.mean_
        cov = (Xc.T @ Xc) / (n - 1)
        evals, evecs = torch.linalg.eigh(cov)
        evals, evecs = evals.flip(0).clamp(min=0.0), evecs.flip(1)
        Vt = evecs.T
        # sign convention of sklearn (svd_flip, u_based_decision=False): the entry of largest
        # magnitude of every component is positive
        idx = Vt.abs().argmax(dim=1)
        Vt = Vt * torch.sign(Vt[torch.arange(Vt.shape[0], device=Vt.device), idx])[:, None]
        self.components_ = Vt[:self.n_components].contiguous()
        self.explained_variance_ = evals[:self.n_components]
        self.explained_variance_ratio_ = self.explained_variance_ / evals.sum()
        return self

    def transform(self, X):
        return (X.to(torch.float64) - self.mean_) @ self.components_.T


def initial_latents(*, Y, mask, v, h, latent_dim, num_samples=100000, whiten=True, pca=None, Gamma=None,
                    sample_index=None, key=0, **kwargs):
    """initialize.py:20-37 -> (latents (N,T,latent_dim) float32, Cd ((k-1) d, latent_dim+1) float64, pca).
    Centred, egocentrically aligned keypoints are projected on the leading principal components of
    a sample of `num_samples` frames (drawn with replacement; `sample_index` injects the indices),
    then whitened with the Cholesky factor of the latents' covariance over unmasked frames."""
    Y = torch.as_tensor(Y)
    dev = Y.device
    n, t, k, d = Y.shape
    Gamma = center_embedding(k, dev) if Gamma is None else f64(Gamma, dev)
    ego = inverse_affine_transform(Y.to(torch.float64), f64(v, dev), f64(h, dev))
    y = torch.einsum("ji,ntjd->ntid", Gamma, ego)                      # Gamma^T applied over keypoints
    yflat = y.reshape(n * t, (k - 1) * d)
    if pca is None:
        if sample_index is None:
            gen = dist.torch_generator(dist.seed_of(key), 21, dev)
            sample_index = torch.randint(0, n * t, (int(num_samples),), device=dev, generator=gen)
        pca = PCAModel(latent_dim).fit(yflat[torch.as_tensor(sample_index, device=dev).long()])
    latents = pca.transform(yflat)[:, :latent_dim]
    comps = torch.as_tensor(pca.components_, device=dev).to(torch.float64)
    mean = torch.as_tensor(pca.mean_, device=dev).to(torch.float64)
    Cd = torch.cat([comps.T, mean[:, None]], dim=1)
    if whiten:
        keep = latents[torch.as_tensor(mask, device=dev).reshape(-1) > 0]
        cov = torch.cov(keep.T)
        L = torch.linalg.cholesky(cov)
        latents = torch.linalg.solve_triangular(L, latents.T, upper=False).T      # latents @ inv(L)^T
        Cd = torch.cat([Cd[:, :-1] @ L, Cd[:, -1:]], dim=1)
    return latents.reshape(n, t, latent_dim).to(torch.float32).contiguous(), Cd.contiguous(), pca


def initial_location(*, Y, **kwargs):
    """initialize.py:39-40: centroid of the keypoints, zero beyond the first two coordinates."""
    Y = torch.as_tensor(Y)
    v = Y.mean(-2)
    v[..., 2:] = 0
    return v


def initial_heading(posterior_keypoints, anterior_keypoints, *, Y, **kwargs):
    """initialize.py:42-45: direction from the mean posterior to the mean anterior keypoint."""
    Y = torch.as_tensor(Y)
    post = torch.as_tensor(posterior_keypoints, device=Y.device).long()
    ante = torch.as_tensor(anterior_keypoints, device=Y.device).long()
    return vector_to_angle(Y[..., ante, :2].mean(-2) - Y[..., post, :2].mean(-2))


def initial_ar_params(key, *, num_states, nu_0, S_0, M_0, K_0, draws=None, device=None, **kwargs):
    """initialize.py:48-51: num_states independent draws of (Ab, Q) from the MNIW prior."""
    S_0 = f64(S_0, device)
    dev = S_0.device
    M_0, K_0 = f64(M_0, dev), f64(K_0, dev)
    K, D, P1 = int(num_states), S_0.shape[-1], M_0.shape[-1]
    if draws is not None:
        chi2, normals, Gd = f64(draws["chi2"], dev), f64(draws["normals"], dev), f64(draws["G"], dev)
    else:
        gen = dist.torch_generator(dist.seed_of(key), 15, dev)
        dof = float(nu_0) - torch.arange(D, dtype=torch.float64, device=dev)
        chi2 = dist.sample_gamma(gen, (dof / 2).expand(K, D).contiguous()) * 2
        normals = torch.randn((K, D * (D - 1) // 2), dtype=torch.float64, device=dev, generator=gen)
        Gd = torch.randn((K, D, P1), dtype=torch.float64, device=dev, generator=gen)
    nu = torch.full((K,), float(nu_0), dtype=torch.float64, device=dev)
    Ab, Q = dist.sample_mniw(nu, S_0.expand(K, D, D), M_0.expand(K, D, P1), K_0.expand(K, P1, P1), chi2, normals, Gd)
    return Ab.contiguous(), Q.contiguous()


def initial_hdp_transitions(key, *, num_states, alpha, kappa, gamma, draws=None, device="cuda", **kwargs):
    """initialize.py:53-58 -> (betas, pi): betas ~ Dir(gamma/K), then one hdp step on zero counts."""
    K = int(num_states)
    counts = torch.zeros((K, K), dtype=torch.int64, device=device)
    if draws is not None and "betas_init_g" in draws:
        g = f64(draws["betas_init_g"], device)
        betas_init = g / g.sum()
    else:
        gen = dist.torch_generator(dist.seed_of(key), 16, device)
        betas_init = dist.sample_dirichlet(gen, torch.full((K,), float(gamma) / K, dtype=torch.float64, device=device))
    return G.sample_hdp_transitions(key, counts, betas_init, alpha, kappa, gamma, draws=draws)


def initial_transitions(key, *, num_states, alpha, kappa, draws=None, device="cuda", **kwargs):
    """initialize.py:60-64 -> pi from the sticky Dirichlet prior (zero counts)."""
    K = int(num_states)
    counts = torch.zeros((K, K), dtype=torch.int64, device=device)
    return G.sample_transitions(key, counts, alpha, kappa, draws=draws)
