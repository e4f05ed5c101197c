"""CPU oracle for the keypoint-SLDS blocked Gibbs sweep.  TEST INFRASTRUCTURE ONLY.

This module is a float64 NumPy (+ numba for the sequential loops) restatement of
the reference algorithm (calebweinreb/keypointMoSeq, `keypoint_moseq/*.py`).
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` /
`--impl reference` legs may import it.  The product (`keypointmoseq_b200`)
never imports it and has no CPU fallback.

PARITY STATUS: pinned against the reference's own source, with the caveat below.  The
reference ships no tests, no golden vectors and its example data is stripped, and JAX /
tensorflow-probability are not installable in this image, so the reference cannot run as
shipped.  Instead (tests/test_reference_pins.py, fixtures under tests/golden/):

(1) its two numba functions on the path (`transitions.py:7-21` sample_crp_tablecounts with
    numba's seeded Mersenne-Twister stream replayed as injected uniforms, `util.py:178-189`
    count_transitions) were executed unmodified: the oracle and the CUDA kernels equal their
    outputs bit for bit;
(2) the unmodified files gibbs.py, kalman.py, distributions.py and util.py were executed under
    a NumPy stand-in for the JAX calls they make (`tests/golden/jax_numpy_shim.py`: jit =
    identity, vmap / lax.scan / lax.map / lax.cond = Python loops, jnp = numpy, random draws
    injected): every resample_* step except resample_hdp_transitions; the oracle agrees with
    those runs to 1e-12 .. 1e-16 (state sequences identical) and the CUDA path to its stated
    float32 tolerances.

Caveat: what is NOT the reference's code in (2) are the third-party samplers and densities it
delegates to -- jr.categorical, jr.gamma, jr.normal / multivariate_normal, tfd.VonMises.sample,
tfd.MultivariateNormalFullCovariance.log_prob -- restated as described below; XLA's own floating
point behaviour is likewise not reproduced.  In addition the independent known-answer tests in
`tests/test_oracle_known_answers.py` (brute-force HMM enumeration, dense joint-Gaussian Kalman
posterior, naive-loop statistics) check the algorithms themselves.

Every random quantity is INJECTED (no RNG in here) so that the CUDA path and the
oracle consume identical draws:

* categorical state draws: one uniform per frame, inverse-CDF in state order
  (`z = first j with cumsum(w)_j > u * sum(w)`); the reference uses
  `jr.categorical` = Gumbel-max (`distributions.py:20-21,65`), which cannot be
  replayed without JAX's threefry stream, so inverse-CDF is the graded contract.
* Gaussian draws: standard normals `w`; `x = chol_lower(S) @ w + m`
  (`kalman.py:139-140,150`).
* gamma / chi2 draws: already distributed with the shape the call site needs.
* von Mises: Best-Fisher rejection driven by injected uniforms (R rounds).

All citations are relative to /root/reference.
"""
import math
import numpy as np

try:  # numba only accelerates the sequential scans; semantics are unchanged
    from numba import njit, prange
    HAVE_NUMBA = True
except Exception:  # pragma: no cover
    HAVE_NUMBA = False

    def njit(*a, **k):
        if len(a) == 1 and callable(a[0]) and not k:
            return a[0]
        return lambda f: f
    prange = range


# ----------------------------------------------------------------------------
# util.py
# ----------------------------------------------------------------------------

def center_embedding(k):
    """util.py:38-39.  First k-1 left singular vectors of I - 11^T/k.  The basis
    of the (k-1)-dim subspace is arbitrary (all singular values are 1), so every
    caller passes the resulting Gamma explicitly to both oracle and CUDA path."""
    return np.linalg.svd(np.eye(k) - np.ones((k, k)) / k)[0][:, :-1]


def pad_affine(x):
    """util.py:169-176."""
    return np.concatenate([x, np.ones((*x.shape[:-1], 1))], axis=-1)


def get_lags(x, nlags):
    """util.py:192-211.  Row t of the result is [x[t-nlags], ..., x[t-1]]
    (oldest lag first) for t = nlags..T-1."""
    lags = [np.roll(x, t, axis=-2) for t in range(1, nlags + 1)]
    return np.concatenate(lags[::-1], axis=-1)[..., nlags:, :]


def ar_to_lds(As, bs, Qs, Cs):
    """util.py:214-230.  Companion form; 1e-2 jitter on the lag copies is part of
    the model."""
    k, l = As.shape[1], As.shape[2] // As.shape[1]
    A_ = np.zeros((As.shape[0], k * l, k * l))
    A_[:, :-k, k:] = np.eye(k * (l - 1))
    A_[:, -k:] = As
    Q_ = np.zeros((Qs.shape[0], k * l, k * l))
    Q_[:, :-k, :-k] = np.eye(k * (l - 1)) * 1e-2
    Q_[:, -k:, -k:] = Qs
    b_ = np.zeros((bs.shape[0], k * l))
    b_[:, -k:] = bs
    C_ = np.zeros((Cs.shape[0], k * l))
    C_[:, -k:] = Cs
    return A_, b_, Q_, C_


def vector_to_angle(V):
    """util.py:66-82 (the 1e-10 offset is mirrored)."""
    y, x = V[..., 1], V[..., 0] + 1e-10
    return (np.arctan(y / x) + (x > 0) * np.pi) % (2 * np.pi) - np.pi


def angle_to_rotation_matrix(h, keypoint_dim=3):
    """util.py:85-109."""
    m = np.tile(np.eye(keypoint_dim), (*h.shape, 1, 1))
    m[..., 0, 0] = np.cos(h)
    m[..., 1, 1] = np.cos(h)
    m[..., 0, 1] = -np.sin(h)
    m[..., 1, 0] = np.sin(h)
    return m


def affine_transform(Y, v, h):
    """util.py:112-138.  Y -> R(h) Y + v."""
    rot = angle_to_rotation_matrix(h, keypoint_dim=Y.shape[-1])
    return (Y[..., None, :] * rot[..., None, :, :]).sum(-1) + v[..., None, :]


def inverse_affine_transform(Y, v, h):
    """util.py:140-166.  Y -> R(-h) (Y - v)."""
    rot = angle_to_rotation_matrix(-h, keypoint_dim=Y.shape[-1])
    return ((Y - v[..., None, :])[..., None, :] * rot[..., None, :, :]).sum(-1)


def count_transitions(num_states, mask, stateseqs):
    """util.py:178-189 (serial numba loop in the reference)."""
    counts = np.zeros((num_states, num_states))
    ok = (mask[:, :-1] > 0) & (mask[:, 1:] > 0)
    np.add.at(counts, (stateseqs[:, :-1][ok], stateseqs[:, 1:][ok]), 1.0)
    return counts


def merge_data(data_dict, keys=None, batch_length=None):
    """util.py:47-60."""
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


def expected_Ybar(x, Cd, Gamma, k, d):
    """The `Ybar = Gamma @ (pad_affine(x)@Cd.T).reshape(...,k-1,d)` idiom of
    gibbs.py:29,43,92,121."""
    return Gamma @ (pad_affine(x) @ Cd.T).reshape(*x.shape[:-1], k - 1, d)


# ----------------------------------------------------------------------------
# distributions.py : HMM forward filter / backward sampler
# ----------------------------------------------------------------------------

@njit(cache=True)
def _hmm_forward(ll, mask, pi):
    """distributions.py:51-60,74-75.  Scaled forward filter, uniform initial
    distribution, mask-blended carry, +1e-16 guard."""
    T, K = ll.shape
    alphan = np.empty((T, K))
    in_pot = np.full(K, 1.0 / K)
    logtot = 0.0
    for t in range(T):
        cmax = ll[t].max()
        a = in_pot * np.exp(ll[t] - cmax)
        norm = a.sum() + 1e-16
        a = a / norm
        logprob = np.log(norm) + cmax
        m = mask[t]
        in_pot = np.dot(a, pi) * m + in_pot * (1.0 - m)
        logtot += logprob * m
        alphan[t] = a
    return alphan, logtot


@njit(cache=True)
def _hmm_backward(alphan, mask, pi, u):
    """distributions.py:62-72,77-78 with the categorical draw replaced by
    inverse-CDF on one injected uniform per frame (see module docstring)."""
    T, K = alphan.shape
    z = np.zeros(T, dtype=np.int64)
    next_pot = np.ones(K)
    for t in range(T - 1, -1, -1):
        if mask[t] > 0:
            c = np.cumsum(next_pot * alphan[t])
            thr = u[t] * c[K - 1]
            s = K - 1
            for j in range(K):
                if c[j] > thr:
                    s = j
                    break
            z[t] = s
            next_pot = pi[:, s].copy()
        else:
            z[t] = 0
    return z


def sample_hmm_stateseq(ll, mask, pi, u):
    """distributions.py:46-79 -> (stateseq, marginal log-likelihood)."""
    alphan, loglik = _hmm_forward(np.ascontiguousarray(ll, dtype=np.float64),
                                  np.ascontiguousarray(mask, dtype=np.float64),
                                  np.ascontiguousarray(pi, dtype=np.float64))
    z = _hmm_backward(alphan, np.ascontiguousarray(mask, dtype=np.float64),
                      np.ascontiguousarray(pi, dtype=np.float64),
                      np.ascontiguousarray(u, dtype=np.float64))
    return z, loglik


@njit(cache=True, parallel=True)
def _hmm_batch(ll, mask, pi, u):
    N, T, K = ll.shape
    z = np.zeros((N, T), dtype=np.int64)
    loglik = np.zeros(N)
    for n in prange(N):
        alphan, lt = _hmm_forward(ll[n], mask[n], pi)
        z[n] = _hmm_backward(alphan, mask[n], pi, u[n])
        loglik[n] = lt
    return z, loglik


def mvn_full_log_prob(x, mu, Q):
    """tfd.MultivariateNormalFullCovariance(mu, Q).log_prob(x) (third-party;
    call site gibbs.py:104): -1/2 |L^-1 (x-mu)|^2 - sum log diag L - D/2 log 2pi."""
    import scipy.linalg as sla
    L = np.linalg.cholesky(Q)
    r = sla.solve_triangular(L, (x - mu).reshape(-1, x.shape[-1]).T, lower=True)
    out = -0.5 * (r ** 2).sum(0) - np.log(np.diag(L)).sum() - 0.5 * x.shape[-1] * np.log(2 * np.pi)
    return out.reshape(x.shape[:-1])


def ar_log_likelihood(x, Ab, Q):
    """gibbs.py:100-104,110: (N, T-L, K) emission log-likelihoods (after the
    reference's moveaxis)."""
    K = Ab.shape[0]
    nlags = Ab.shape[2] // Ab.shape[1]
    xin = pad_affine(get_lags(x, nlags))
    out = np.empty((*xin.shape[:-1], K))
    for k in range(K):
        mu = xin @ Ab[k].T
        out[..., k] = mvn_full_log_prob(x[..., nlags:, :], mu, Q[k])
    return out


def resample_stateseqs(*, x, mask, Ab, Q, pi, u, **kwargs):
    """gibbs.py:107-115 -> (z (N,T-L) int64, loglik (N,)).  `u`: (N,T-L)."""
    nlags = Ab.shape[2] // Ab.shape[1]
    ll = ar_log_likelihood(x, Ab, Q)
    m = np.ascontiguousarray(mask[:, nlags:], dtype=np.float64)
    return _hmm_batch(np.ascontiguousarray(ll), m, np.ascontiguousarray(pi, dtype=np.float64),
                      np.ascontiguousarray(u, dtype=np.float64))


# ----------------------------------------------------------------------------
# kalman.py (live code only, :80-156)
# ----------------------------------------------------------------------------

@njit(cache=True)
def _condition_on(m, S, C, D, R, y):
    """kalman.py:91-95."""
    Sinv = np.linalg.inv(S)
    CtR = C.T / R
    S_cond = np.linalg.inv(Sinv + CtR @ C)
    m_cond = S_cond @ (Sinv @ m + CtR @ (y - D))
    return m_cond, S_cond


@njit(cache=True)
def _obs_logpdf(m, S, C, D, R, y):
    """Innovation-form log N(y; C m + D, C S C^T + diag R).  NOT in the reference
    (kalman.py:81-84 promises a marginal likelihood, the code computes none); it
    is defined here because BASELINE.json grades a Kalman marginal log-lik."""
    r = y - D - C @ m
    V = C @ S @ C.T + np.diag(R)
    L = np.linalg.cholesky(V)
    a = np.linalg.solve(L, r)
    return -0.5 * (a * a).sum() - np.log(np.diag(L)).sum() - 0.5 * y.shape[0] * np.log(2 * np.pi)


@njit(cache=True)
def kalman_filter(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs):
    """kalman.py:80-120.  Returns (filtered_ms (T',n), filtered_Ss (T',n,n),
    loglik).  Masked steps emit the *predicted* moments and leave the carry
    unchanged (:109-111); the final update is unconditional (:117)."""
    Tp = ys.shape[0]
    n = m0.shape[0]
    ms = np.empty((Tp, n))
    Ss = np.empty((Tp, n, n))
    m_pred = m0.copy()
    S_pred = S0.copy()
    ll = 0.0
    for t in range(Tp - 1):
        if mask[t] > 0:
            ll += _obs_logpdf(m_pred, S_pred, C, D, Rs[t], ys[t])
            m_c, S_c = _condition_on(m_pred, S_pred, C, D, Rs[t], ys[t])
            ms[t] = m_c
            Ss[t] = S_c
            z = zs[t]
            m_pred = A[z] @ m_c + B[z]                  # :86-89
            S_pred = A[z] @ S_c @ A[z].T + Q[z]
        else:
            ms[t] = m_pred
            Ss[t] = S_pred
    ll += _obs_logpdf(m_pred, S_pred, C, D, Rs[Tp - 1], ys[Tp - 1])
    m_c, S_c = _condition_on(m_pred, S_pred, C, D, Rs[Tp - 1], ys[Tp - 1])
    ms[Tp - 1] = m_c
    Ss[Tp - 1] = S_c
    return ms, Ss, ll


@njit(cache=True)
def kalman_sample(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, w, w_last):
    """kalman.py:123-156.  `w` (T'-1,n) replaces jr.normal (:147); `w_last` (n)
    drives the multivariate-normal draw of the last state (:150, Cholesky
    method: m + chol(S) w).  Returns (xs (T',n), loglik)."""
    ms, Ss, ll = kalman_filter(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs)
    Tp = ys.shape[0]
    n = m0.shape[0]
    K = Q.shape[0]
    Qinv = np.empty_like(Q)
    for k in range(K):
        Qinv[k] = np.linalg.inv(Q[k])                    # :146
    xs = np.zeros((Tp, n))
    x = ms[Tp - 1] + np.linalg.cholesky(Ss[Tp - 1]) @ w_last
    xs[Tp - 1] = x
    for t in range(Tp - 2, -1, -1):
        if mask[t] > 0:
            z = zs[t]
            Sinv = np.linalg.inv(Ss[t])                  # :129-133
            AtQ = A[z].T @ Qinv[z]
            S_cond = np.linalg.inv(Sinv + AtQ @ A[z])
            m_cond = S_cond @ (Sinv @ ms[t] + AtQ @ (x - B[z]))
            L = np.linalg.cholesky(S_cond)               # :138-140
            x = L @ w[t] + m_cond
            xs[t] = x
        # masked: emit zeros, carry x unchanged (:142-143)
    return xs, ll


@njit(cache=True, parallel=True)
def _kalman_sample_batch(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, w, w_last):
    N, Tp, p = ys.shape
    n = m0.shape[0]
    out = np.zeros((N, Tp, n))
    lls = np.zeros(N)
    for i in prange(N):
        xs, ll = kalman_sample(ys[i], mask[i], zs[i], m0, S0, A, B, Q, C, D, Rs[i], w[i], w_last[i])
        out[i] = xs
        lls[i] = ll
    return out, lls


def _c(a, dt=np.float64):
    return np.ascontiguousarray(a, dtype=dt)


# ----------------------------------------------------------------------------
# gibbs.py
# ----------------------------------------------------------------------------

def resample_latents(*, Y, mask, v, h, z, s, Cd, sigmasq, Ab, Q, Gamma, w, w_last,
                     return_loglik=False, **kwargs):
    """gibbs.py:10-23.  `w`: (N, T-L, n), `w_last`: (N, n) with n = D*L."""
    d, nlags, n = Ab.shape[1], Ab.shape[2] // Ab.shape[1], Y.shape[0]
    Cdf = np.kron(Gamma, np.eye(Y.shape[-1])) @ Cd                       # :14
    ys = inverse_affine_transform(Y, v, h).reshape(*Y.shape[:-2], -1)    # :15
    A, B, Q_, C = ar_to_lds(Ab[..., :-1], Ab[..., -1], Q, Cdf[..., :-1])  # :16
    D = Cdf[..., -1]
    R = np.repeat(s * sigmasq, Y.shape[-1], axis=-1)[:, nlags - 1:]      # :17
    mu0, S0 = np.zeros(d * nlags), np.eye(d * nlags) * 10                # :18
    xs, lls = _kalman_sample_batch(
        _c(ys[:, nlags - 1:]), _c(mask[:, nlags - 1:-1]), _c(z, np.int64), mu0, S0,
        _c(A), _c(B), _c(Q_), _c(C), _c(D), _c(R), _c(w), _c(w_last))
    xs = np.concatenate([xs[:, 0, :-d].reshape(-1, nlags - 1, d), xs[:, :, -d:]], axis=1)  # :22
    return (xs, lls) if return_loglik else xs


def heading_stats(*, Y, v, x, s, Cd, sigmasq, Gamma, **kwargs):
    """gibbs.py:27-34 -> (theta, kappa) of the von Mises conditional."""
    k, d = Y.shape[-2:]
    Ybar = expected_Ybar(x, Cd, Gamma, k, d)
    S = (Ybar[..., None] * (Y - v[..., None, :])[..., None, :] / (s * sigmasq)[..., None, None]).sum(-3)
    kappa_cos = S[..., 0, 0] + S[..., 1, 1]
    kappa_sin = S[..., 0, 1] - S[..., 1, 0]
    theta = vector_to_angle(np.stack([kappa_cos, kappa_sin], axis=-1))
    kappa = np.sqrt(kappa_cos ** 2 + kappa_sin ** 2)
    return theta, kappa


def sample_vonmises(theta, kappa, u):
    """distributions.py:5-6 delegates to tfd.VonMises(theta,kappa).sample (a
    rejection sampler whose draw stream cannot be replayed).  Contract used by
    both the oracle and the CUDA path: Best & Fisher (1979) rejection with
    injected uniforms u[..., r, 0:3] for round r; the first accepted round wins;
    if no round accepts, the mean direction is returned.  kappa < 1e-8 is
    treated as uniform on the circle.  Result wrapped to [-pi, pi]."""
    theta = np.asarray(theta, dtype=np.float64)
    kappa = np.asarray(kappa, dtype=np.float64)
    out = np.array(theta, copy=True)
    it = np.nditer(theta, flags=['multi_index'])
    for _ in it:
        idx = it.multi_index
        kap = float(kappa[idx])
        uu = u[idx]
        dev = 0.0
        if kap < 1e-8:
            dev = math.pi * (2.0 * uu[0, 2] - 1.0)
        else:
            tau = 1.0 + math.sqrt(1.0 + 4.0 * kap * kap)
            rho = (tau - math.sqrt(2.0 * tau)) / (2.0 * kap)
            r = (1.0 + rho * rho) / (2.0 * rho)
            for rnd in range(uu.shape[0]):
                u1, u2, u3 = uu[rnd]
                zc = math.cos(math.pi * u1)
                f = (1.0 + r * zc) / (r + zc)
                c = kap * (r - f)
                if c * (2.0 - c) - u2 > 0.0 or math.log(c / u2) + 1.0 - c >= 0.0:
                    f = min(1.0, max(-1.0, f))
                    dev = math.acos(f) if u3 > 0.5 else -math.acos(f)
                    break
        hh = float(theta[idx]) + dev
        out[idx] = hh - 2.0 * math.pi * round(hh / (2.0 * math.pi))
    return out


def resample_heading(*, Y, v, x, s, Cd, sigmasq, Gamma, u, **kwargs):
    """gibbs.py:25-35.  `u`: (N,T,R,3) uniforms for the von Mises sampler."""
    theta, kappa = heading_stats(Y=Y, v=v, x=x, s=s, Cd=Cd, sigmasq=sigmasq, Gamma=Gamma)
    return sample_vonmises(theta, kappa, u)


def location_stats(*, Y, h, x, s, Cd, sigmasq, Gamma, **kwargs):
    """gibbs.py:40-46 -> (mu (N,T,d), gammasq (N,T))."""
    k, d = Y.shape[-2:]
    gammasq = 1 / (1 / (s * sigmasq)).sum(-1)
    Ybar = expected_Ybar(x, Cd, Gamma, k, d)
    rot = angle_to_rotation_matrix(h, keypoint_dim=d)
    mu = ((Y - (rot[..., None, :, :] * Ybar[..., None, :]).sum(-1))
          * (gammasq[..., None] / (s * sigmasq))[..., None]).sum(-2)
    return mu, gammasq


def resample_location(*, mask, Y, h, x, s, Cd, sigmasq, sigmasq_loc, Gamma, w, w_last,
                      return_loglik=False, **kwargs):
    """gibbs.py:38-54.  `w`: (N,T-1,d), `w_last`: (N,d)."""
    k, d = Y.shape[-2:]
    mu, gammasq = location_stats(Y=Y, h=h, x=x, s=s, Cd=Cd, sigmasq=sigmasq, Gamma=Gamma)
    m0, S0 = np.zeros(d), np.eye(d) * 1e6
    A, B, Q = np.eye(d)[None], np.zeros(d)[None], np.eye(d)[None] * sigmasq_loc
    C, D, R = np.eye(d), np.zeros(d), gammasq[..., None] * np.ones(d)
    z = np.zeros(mask[:, 1:].shape, dtype=np.int64)
    xs, lls = _kalman_sample_batch(_c(mu), _c(mask[:, :-1]), z, m0, S0, _c(A), _c(B), _c(Q),
                                   _c(C), _c(D), _c(R), _c(w), _c(w_last))
    return (xs, lls) if return_loglik else xs


def obs_variance_stats(*, Y, mask, Cd, v, h, x, s, Gamma, **kwargs):
    """gibbs.py:86-94 -> (S_y (k,), mask.sum())."""
    k, d = Y.shape[-2:]
    s_ = s.reshape(-1, k)
    m = mask.flatten()
    x_ = x.reshape(-1, x.shape[-1])
    Ybar = expected_Ybar(x_, Cd, Gamma, k, d)
    Ytild = inverse_affine_transform(Y, v, h).reshape(-1, k, d)
    S_y = (m[:, None] * ((Ytild - Ybar) ** 2).sum(-1) / s_).sum(0)
    return S_y, m.sum()


def resample_obs_variance(*, Y, mask, Cd, v, h, x, s, nu_sigma, sigmasq_0, Gamma, g, **kwargs):
    """gibbs.py:84-96.  `g` (k,) ~ Gamma(degs/2, 1) with degs = nu_sigma + 3*mask.sum()
    (the literal 3 is the reference's, independent of the keypoint dim)."""
    S_y, msum = obs_variance_stats(Y=Y, mask=mask, Cd=Cd, v=v, h=h, x=x, s=s, Gamma=Gamma)
    variance = (nu_sigma * sigmasq_0 + S_y) / (nu_sigma + 3 * msum)
    degs = (nu_sigma + 3 * msum) * np.ones_like(variance)
    return sample_scaled_inv_chi2(degs, variance, g)


def obs_params_stats(*, Y, mask, sigmasq, v, h, x, s, **kwargs):
    """The per-keypoint weighted moments behind gibbs.py:58-82: with w_fj = mask_f / (s_fj sigmasq_j),
    W[j] = sum_f w_fj xt_f xt_f^T  (D+1, D+1)  and  V[j] = sum_f w_fj xt_f Ytild_fj^T  (D+1, d),
    xt = pad_affine(x).  (S_xx = sum_j Gamma_j Gamma_j^T (x) W[j], S_yx = sum_j Gamma_j (x) V[j].)"""
    k, d = Y.shape[-2:]
    D = x.shape[-1]
    m = mask.reshape(-1)
    sf = s.reshape(-1, k)
    xt = pad_affine(x.reshape(-1, D))
    Yt = inverse_affine_transform(Y, v, h).reshape(-1, k, d)
    w = m[:, None] / (sf * sigmasq[None, :])                       # (F, k)
    W = np.einsum('fj,fa,fb->jab', w, xt, xt)
    V = np.einsum('fj,fa,fjc->jac', w, xt, Yt)
    return W, V


def obs_params_posterior(W, V, Gamma, sigmasq_C):
    """gibbs.py:69-80 from the moments: S_xx[(a,i),(b,l)] (x) I_d, S_yx[((a,i),c)], then
    Sigma_n = inv(I / sigmasq_C + S_xx), mu_n = Sigma_n S_yx."""
    k, D1, d = V.shape
    S4 = np.einsum('ji,jl,jab->aibl', Gamma, Gamma, W)                # (D+1, k-1, D+1, k-1)
    S_xx = np.kron(S4.reshape(D1 * (k - 1), D1 * (k - 1)), np.eye(d))
    S_yx = np.einsum('ji,jac->aic', Gamma, V).reshape(-1)
    Sigma_n = np.linalg.inv(np.eye(d * D1 * (k - 1)) / sigmasq_C + S_xx)
    return Sigma_n @ S_yx, Sigma_n


def resample_obs_params(*, Y, mask, sigmasq, v, h, x, s, sigmasq_C, Gamma, w, **kwargs):
    """gibbs.py:58-82 -> Cd ((k-1) d, D+1).  `w` (d (D+1) (k-1),) standard normals for
    jr.multivariate_normal(key, mu_n, Sigma_n) = mu_n + cholesky(Sigma_n) @ w."""
    k, d = Y.shape[-2:]
    D = x.shape[-1]
    W, V = obs_params_stats(Y=Y, mask=mask, sigmasq=sigmasq, v=v, h=h, x=x, s=s)
    mu_n, Sigma_n = obs_params_posterior(W, V, Gamma, sigmasq_C)
    draw = mu_n + np.linalg.cholesky(Sigma_n) @ w
    return draw.reshape(D + 1, d * (k - 1)).T


def resample_scales(*, x, v, h, Y, Cd, sigmasq, nu_s, s_0, Gamma, g, **kwargs):
    """gibbs.py:117-125.  `g` (N,T,k) ~ Gamma((nu_s+3)/2, 1).  No mask."""
    k, d = Y.shape[-2:]
    Ybar = expected_Ybar(x, Cd, Gamma, k, d)
    Ytild = inverse_affine_transform(Y, v, h)
    variance = (((Ytild - Ybar) ** 2).sum(-1) / sigmasq + s_0 * nu_s) / (nu_s + 3)
    degs = (nu_s + 3) * np.ones_like(variance)
    return sample_scaled_inv_chi2(degs, variance, g)


def sample_scaled_inv_chi2(degs, variance, g):
    """distributions.py:8-15: 1 / (Gamma(degs/2) / (degs*variance/2))."""
    return 1.0 / (g / (degs * variance / 2))


# ---- AR parameters -----------------------------------------------------------

def ar_sufficient_stats(*, nlags, num_states, mask, x, z, **kwargs):
    """The masked second moments of gibbs.py:129-131,141-143, returned as one
    symmetric (K, P, P) matrix over f = [x_in (D*L), 1, x_out (D)], P = D*L+1+D:
    S_in_in = F[:, :DL+1, :DL+1], S_out_in = F[:, DL+1:, :DL+1],
    S_out_out = F[:, DL+1:, DL+1:], count = F[:, DL, DL]."""
    D = x.shape[-1]
    x_in = pad_affine(get_lags(x, nlags)).reshape(-1, nlags * D + 1)
    x_out = x[..., nlags:, :].reshape(-1, D)
    f = np.concatenate([x_in, x_out], axis=1)
    m = mask[..., nlags:].reshape(-1)
    zf = z.reshape(-1)
    F = np.zeros((num_states, f.shape[1], f.shape[1]))
    for k in range(num_states):
        sel = (zf == k) & (m > 0)
        fk = f[sel] * m[sel][:, None] ** 0.5
        F[k] = fk.T @ fk
    return F


def sample_invwishart(S, nu, chi2, normals):
    """distributions.py:29-37.  `chi2` (n,) ~ chi2(nu - arange(n)); `normals`
    (n(n-1)/2,) fill triu_indices(n,1) in row-major order."""
    import scipy.linalg as sla
    n = S.shape[0]
    chol = np.linalg.cholesky(S)
    xm = np.diag(np.sqrt(chi2))
    xm[np.triu_indices(n, 1)] = normals
    R = np.linalg.qr(xm, mode='r')
    T = sla.solve_triangular(R.T, chol.T, lower=True).T
    return T @ T.T


def sample_mn(M, U, V, G):
    """distributions.py:23-27."""
    G = np.linalg.cholesky(U) @ G
    G = G @ np.linalg.cholesky(V).T
    return M + G


def sample_mniw(nu, S, M, K, chi2, normals, G):
    """distributions.py:39-42."""
    sigma = sample_invwishart(S, nu, chi2, normals)
    A = sample_mn(M, sigma, K, G)
    return A, sigma


def regression_posterior(S_out_out, S_out_in, S_in_in, nu_0, S_0, M_0, K_0, count):
    """gibbs.py:132-136 -> (nu_n, S_n, M_n, K_n)."""
    K_0_inv = np.linalg.inv(K_0)
    K_n = np.linalg.inv(K_0_inv + S_in_in)
    M_n = (M_0 @ K_0_inv + S_out_in) @ K_n
    S_n = S_0 + S_out_out + (M_0 @ K_0_inv @ M_0.T - M_n @ np.linalg.inv(K_n) @ M_n.T)
    return nu_0 + count, S_n, M_n, K_n


def resample_ar_params(*, nlags, num_states, mask, x, z, nu_0, S_0, M_0, K_0,
                       chi2, normals, G, **kwargs):
    """gibbs.py:139-145 (+ :127-136).  `chi2` (K,D), `normals` (K,D(D-1)/2),
    `G` (K,D,D*L+1)."""
    D = x.shape[-1]
    P1 = nlags * D + 1
    F = ar_sufficient_stats(nlags=nlags, num_states=num_states, mask=mask, x=x, z=z)
    Ab = np.empty((num_states, D, P1))
    Q = np.empty((num_states, D, D))
    for k in range(num_states):
        nu_n, S_n, M_n, K_n = regression_posterior(
            F[k, P1:, P1:], F[k, P1:, :P1], F[k, :P1, :P1], nu_0, S_0, M_0, K_0, F[k, P1 - 1, P1 - 1])
        Ab[k], Q[k] = sample_mniw(nu_n, S_n, M_n, K_n, chi2[k], normals[k], G[k])
    return Ab, Q


# ---- diagnostics ------------------------------------

def log_joint_likelihood(*, Y, mask, x, s, v, h, z, pi, Ab, Q, Cd, sigmasq, sigmasq_loc, nu_s, s_0,
                         Gamma, **kwargs):
    """util.py:232-275 -> dict of the five masked log-probability sums 'Y','x','z','v','s'."""
    k, d = Y.shape[-2:]
    nlags = Ab.shape[2] // Ab.shape[1]
    na = None
    Ybar = expected_Ybar(x, Cd, Gamma, k, d)
    sqerr = ((Y - affine_transform(Ybar, v, h)) ** 2).sum(-1)
    lpY = -0.5 * sqerr / s / sigmasq - d / 2 * np.log(2 * s * sigmasq * np.pi)
    Qinv, Qdet = np.linalg.inv(Q), np.linalg.det(Q)
    x_pred = (Ab[z] @ pad_affine(get_lags(x, nlags))[..., na])[..., 0]
    dd = x_pred - x[:, nlags:]
    lpx = (-(dd[..., na, :] * Qinv[z] * dd[..., :, na]).sum((2, 3)) / 2
           - np.log(Qdet[z]) / 2 - np.log(2 * np.pi) * Q.shape[-1] / 2)
    lpz = np.log(pi[z[:, :-1], z[:, 1:]])
    dv = v[:, :-1] - v[:, 1:]
    lpv = -(dv ** 2).sum(-1) / sigmasq_loc / 2 - v.shape[-1] / 2 * np.log(sigmasq_loc * 2 * np.pi)
    lps = -nu_s * s_0 / s / 2 - (1 + nu_s / 2) * np.log(s)
    return {'Y': (lpY * mask[:, :, na]).sum(), 'x': (lpx * mask[:, nlags:]).sum(),
            'z': (lpz * mask[:, nlags + 1:]).sum(), 'v': (lpv * mask[:, 1:]).sum(),
            's': (lps * mask[:, :, na]).sum()}


def stateseq_stats(stateseqs, mask):
    """util.py:41-45 -> (usage = bincount of the unmasked states, durations = run lengths of the
    flattened unmasked sequence)."""
    sq = np.asarray(stateseqs).reshape(-1)[mask[..., -stateseqs.shape[-1]:].reshape(-1) > 0]
    usage = np.bincount(sq)
    change = np.flatnonzero(np.diff(sq)) + 1
    edges = np.concatenate([[0], change, [sq.size]])
    return usage, np.diff(edges)


# ---- transitions.py ------------------------------------------------------------

def sample_crp_tablecounts(concentration, customers, colweights, randseq):
    """transitions.py:7-21 with the numba-private `np.random.random(tot)` stream
    replaced by the injected `randseq` (same `starts[i,j]+k` indexing)."""
    customers = np.asarray(customers, dtype=np.int64)
    m = np.zeros_like(customers)
    flat = customers.ravel()
    tmp = np.empty(flat.size, dtype=np.int64)
    tmp[0] = 0
    tmp[1:] = np.cumsum(flat[:-1])
    starts = tmp.reshape(customers.shape)
    for i in range(customers.shape[0]):
        for j in range(customers.shape[1]):
            c = customers[i, j]
            if c == 0:
                continue
            kk = np.arange(c)
            cw = concentration * colweights[j]
            m[i, j] = (randseq[starts[i, j] + kk] < cw / (kk + cw)).sum()
    return m


def sample_ms(counts, betas, alpha, kappa, randseq, u_thin=None, thin=False):
    """transitions.py:23-33.

    thin=False (default) is what the reference RETURNS: its binomial thinning of the sticky diagonal
    (transitions.py:28-31) is assigned through `newms.flat[::K+1][indices] = ...`, and slicing a
    `flat` iterator yields a copy, so the assignment lands in a temporary and `newms` leaves the
    function equal to the raw CRP table counts (pinned by tests/golden/reference_sweep_bench.npz,
    where the reference's own sample_ms return value is recorded).
    thin=True applies the thinning the code intends (the override correction of the sticky HDP-HMM
    sampler): the host `np.random.binomial(ms_ii, rho_i)` realised as the Bernoulli sum
    `u_thin[i, :ms_ii] < rho_i` over injected uniforms (identical in distribution)."""
    ms = sample_crp_tablecounts(alpha, np.array(counts, dtype=np.int64), np.array(betas), randseq)
    newms = ms.copy()
    if not thin:
        return newms
    K = ms.shape[0]
    for i in range(K):
        rho = betas[i] * alpha / (betas[i] * alpha + kappa)
        newms[i, i] = (u_thin[i, :ms[i, i]] < rho).sum()
    return newms


def hdp_transition_concentrations(counts, ms, betas_new, alpha, kappa, gamma):
    """transitions.py:38-39 Dirichlet parameters: betas ~ Dir(ms.sum(0)+gamma/K),
    pi_i ~ Dir(alpha*betas + counts_i + kappa*e_i)."""
    K = counts.shape[0]
    return ms.sum(0) + gamma / K, alpha * betas_new[None, :] + counts + kappa * np.eye(K)


def dirichlet_from_gammas(g):
    """jr.dirichlet = normalised Gamma(conc,1) variates (third-party)."""
    return g / g.sum(-1, keepdims=True)


def resample_hdp_transitions(*, z, mask, betas, alpha, kappa, gamma, num_states, randseq, betas_g, pi_g,
                             u_thin=None, thin=False):
    """gibbs.py:148-151 -> transitions.py:35-40 with every draw injected: `randseq` for the CRP
    (numba's np.random.random stream), `betas_g` (K,) and `pi_g` (K,K) the Gamma variates behind
    the two jr.dirichlet calls (`u_thin`, `thin`: see sample_ms)."""
    counts = count_transitions(num_states, mask[..., -z.shape[-1]:], z)
    ms = sample_ms(counts, betas, alpha, kappa, randseq, u_thin, thin)
    betas_new = dirichlet_from_gammas(np.asarray(betas_g, dtype=np.float64))
    return betas_new, dirichlet_from_gammas(np.asarray(pi_g, dtype=np.float64)), counts, ms


def sample_transitions(counts, alpha, kappa, pi_g):
    """transitions.py:42-44: pi ~ Dir(counts + alpha + kappa I); returns (pi, concentration)."""
    conc = counts + alpha + kappa * np.eye(counts.shape[0])
    return dirichlet_from_gammas(np.asarray(pi_g, dtype=np.float64)), conc


# ---- initialize.py ------------------------------------------------------------------

def initial_location(Y):
    """initialize.py:39-40."""
    v = Y.mean(-2).copy()
    v[..., 2:] = 0
    return v


def initial_heading(posterior_keypoints, anterior_keypoints, Y):
    """initialize.py:42-45."""
    post = Y[..., posterior_keypoints, :2].mean(-2)
    ante = Y[..., anterior_keypoints, :2].mean(-2)
    return vector_to_angle(ante - post)


def initial_latents(*, Y, mask, v, h, latent_dim, sample_index, Gamma, whiten=True):
    """initialize.py:20-37 with the frame sample injected (`np.random.choice(t*n, num_samples)`)
    and the PCA restated from its definition (sklearn.decomposition.PCA, third-party: principal
    axes of the sample covariance, largest-magnitude entry of each axis positive)."""
    n, t, k, d = Y.shape
    y = np.einsum('ji,ntjd->ntid', Gamma, inverse_affine_transform(Y, v, h))
    yflat = y.reshape(n * t, (k - 1) * d)
    ys = yflat[sample_index]
    mean = ys.mean(0)
    ev, evec = np.linalg.eigh(np.cov(ys.T))
    comps = evec[:, ::-1].T[:latent_dim]
    comps = comps * np.sign(comps[np.arange(latent_dim), np.abs(comps).argmax(1)])[:, None]
    lat = (yflat - mean) @ comps.T
    Cd = np.hstack([comps.T, mean[:, None]])
    if whiten:
        cov = np.cov(lat[mask.reshape(-1) > 0].T)
        Lc = np.linalg.cholesky(cov)
        lat = lat @ np.linalg.inv(Lc).T
        Cd[:, :-1] = Cd[:, :-1] @ Lc
    return lat.reshape(n, t, latent_dim), Cd


def initial_ar_params(*, num_states, nu_0, S_0, M_0, K_0, chi2, normals, G):
    """initialize.py:48-51: one sample_mniw draw from the prior per state."""
    out = [sample_mniw(nu_0, S_0, M_0, K_0, chi2[k], normals[k], G[k]) for k in range(num_states)]
    return np.stack([o[0] for o in out]), np.stack([o[1] for o in out])


# ---- full-sweep convenience used by the CPU baseline ---------------------------

def arhmm_sweep(data, states, params, hyp, draws):
    """One AR-HMM sweep with x fixed: examples/keypoint_slds.ipynb cell 9 and
    examples/depth_ar_hmm.ipynb cell 8 (resample_hdp_transitions, resample_ar_params,
    resample_stateseqs)."""
    L, K, mask = hyp['nlags'], hyp['num_states'], data['mask']
    st, pr = dict(states), dict(params)
    counts = count_transitions(K, mask[..., -st['z'].shape[-1]:], st['z'])
    ms = sample_ms(counts, pr['betas'], hyp['alpha'], hyp['kappa'], draws['crp_u'])
    pr['betas'] = dirichlet_from_gammas(draws['betas_g'](ms.sum(0) + hyp['gamma'] / K))
    pr['pi'] = dirichlet_from_gammas(draws['pi_g'](hyp['alpha'] * pr['betas'][None, :] + counts
                                                   + hyp['kappa'] * np.eye(K)))
    pr['Ab'], pr['Q'] = resample_ar_params(nlags=L, num_states=K, mask=mask, x=st['x'], z=st['z'],
                                           nu_0=hyp['nu_0'], S_0=hyp['S_0'], M_0=hyp['M_0'], K_0=hyp['K_0'],
                                           chi2=draws['ar_chi2'], normals=draws['ar_normals'], G=draws['ar_G'])
    st['z'], _ = resample_stateseqs(x=st['x'], mask=mask, Ab=pr['Ab'], Q=pr['Q'], pi=pr['pi'], u=draws['z_u'])
    return st, pr



def sweep(data, states, params, hyp, draws, Gamma):
    """One blocked Gibbs sweep in the order of examples/keypoint_slds.ipynb cell 12
    (steps 1-8).  `draws` holds every injected variate (see `make_sweep_draws`).
    The Dirichlet / gamma draws of the transition step are supplied as uniforms
    /gammas generated by the caller with the correct shapes."""
    L = hyp['nlags']
    K = hyp['num_states']
    Y, mask = data['Y'], data['mask']
    st, pr = dict(states), dict(params)
    pr['Ab'], pr['Q'] = resample_ar_params(nlags=L, num_states=K, mask=mask, x=st['x'], z=st['z'],
                                           nu_0=hyp['nu_0'], S_0=hyp['S_0'], M_0=hyp['M_0'], K_0=hyp['K_0'],
                                           chi2=draws['ar_chi2'], normals=draws['ar_normals'], G=draws['ar_G'])
    pr['sigmasq'] = resample_obs_variance(Y=Y, mask=mask, Cd=pr['Cd'], v=st['v'], h=st['h'], x=st['x'], s=st['s'],
                                          nu_sigma=hyp['nu_sigma'], sigmasq_0=hyp['sigmasq_0'], Gamma=Gamma,
                                          g=draws['sigmasq_g'])
    counts = count_transitions(K, mask[..., -st['z'].shape[-1]:], st['z'])
    ms = sample_ms(counts, pr['betas'], hyp['alpha'], hyp['kappa'], draws['crp_u'])
    pr['betas'] = dirichlet_from_gammas(draws['betas_g'](ms.sum(0) + hyp['gamma'] / K))
    pr['pi'] = dirichlet_from_gammas(draws['pi_g'](hyp['alpha'] * pr['betas'][None, :] + counts
                                                   + hyp['kappa'] * np.eye(K)))
    st['z'], _ = resample_stateseqs(x=st['x'], mask=mask, Ab=pr['Ab'], Q=pr['Q'], pi=pr['pi'], u=draws['z_u'])
    st['x'] = resample_latents(Y=Y, mask=mask, v=st['v'], h=st['h'], z=st['z'], s=st['s'], Cd=pr['Cd'],
                               sigmasq=pr['sigmasq'], Ab=pr['Ab'], Q=pr['Q'], Gamma=Gamma,
                               w=draws['x_w'], w_last=draws['x_w_last'])
    st['h'] = resample_heading(Y=Y, v=st['v'], x=st['x'], s=st['s'], Cd=pr['Cd'], sigmasq=pr['sigmasq'],
                               Gamma=Gamma, u=draws['h_u'])
    st['v'] = resample_location(mask=mask, Y=Y, h=st['h'], x=st['x'], s=st['s'], Cd=pr['Cd'],
                                sigmasq=pr['sigmasq'], sigmasq_loc=hyp['sigmasq_loc'], Gamma=Gamma,
                                w=draws['v_w'], w_last=draws['v_w_last'])
    st['s'] = resample_scales(x=st['x'], v=st['v'], h=st['h'], Y=Y, Cd=pr['Cd'], sigmasq=pr['sigmasq'],
                              nu_s=hyp['nu_s'], s_0=hyp['s_0'], Gamma=Gamma, g=draws['s_g'])
    return st, pr
