"""Known-answer tests that pin the oracle (the reference ships no tests or golden
vectors — SURVEY.md §4, §8c).  Each test checks the restated algorithm against an
independent brute-force construction of the same mathematical object."""
import itertools

import numpy as np
import pytest

from oracle import kpms_oracle as O


# ------------------------------------------------------------------ HMM ----------

def _enumerate_paths(ll, pi):
    T, K = ll.shape
    paths = list(itertools.product(range(K), repeat=T))
    logp = np.empty(len(paths))
    for i, zp in enumerate(paths):
        lp = -np.log(K) + ll[0, zp[0]]
        for t in range(1, T):
            lp += np.log(pi[zp[t - 1], zp[t]]) + ll[t, zp[t]]
        logp[i] = lp
    return paths, logp


def test_hmm_marginal_loglik_matches_enumeration():
    rng = np.random.default_rng(0)
    K, T = 3, 6
    pi = rng.dirichlet(np.ones(K), size=K)
    ll = rng.standard_normal((T, K)) * 3 - 5
    _, logp = _enumerate_paths(ll, pi)
    want = np.logaddexp.reduce(logp)
    _, got = O.sample_hmm_stateseq(ll, np.ones(T), pi, rng.random(T))
    assert got == pytest.approx(want, rel=1e-12)


def test_hmm_backward_sampler_draws_from_exact_posterior():
    rng = np.random.default_rng(1)
    K, T, S = 3, 5, 200_000
    pi = rng.dirichlet(np.ones(K), size=K)
    ll = rng.standard_normal((T, K)) * 2
    paths, logp = _enumerate_paths(ll, pi)
    post = np.exp(logp - np.logaddexp.reduce(logp))
    u = rng.random((S, T))
    z, _ = O._hmm_batch(np.broadcast_to(ll, (S, T, K)).copy(), np.ones((S, T)), pi, u)
    idx = (z * (K ** np.arange(T - 1, -1, -1))).sum(1)
    emp = np.bincount(idx, minlength=K ** T) / S
    # chi-square style bound: every path frequency within 5 sigma
    sigma = np.sqrt(post * (1 - post) / S)
    assert np.all(np.abs(emp - post) < 5 * sigma + 1e-4)


def test_hmm_trailing_mask_equals_prefix_chain():
    rng = np.random.default_rng(2)
    K, T, Tv = 4, 9, 6
    pi = rng.dirichlet(np.ones(K), size=K)
    ll = rng.standard_normal((T, K))
    mask = (np.arange(T) < Tv).astype(float)
    u = rng.random(T)
    z_full, l_full = O.sample_hmm_stateseq(ll, mask, pi, u)
    z_pre, l_pre = O.sample_hmm_stateseq(ll[:Tv], np.ones(Tv), pi, u[:Tv])
    assert l_full == pytest.approx(l_pre, rel=1e-13)
    assert np.array_equal(z_full[:Tv], z_pre)
    assert np.all(z_full[Tv:] == 0)


def test_hmm_all_masked_row():
    K, T = 3, 4
    pi = np.full((K, K), 1.0 / K)
    z, l = O.sample_hmm_stateseq(np.zeros((T, K)), np.zeros(T), pi, np.full(T, 0.5))
    assert np.all(z == 0) and l == 0.0


def test_ar_loglik_matches_direct_density():
    rng = np.random.default_rng(3)
    N, T, D, L, K = 2, 7, 3, 2, 4
    x = rng.standard_normal((N, T, D))
    Ab = rng.standard_normal((K, D, D * L + 1)) * 0.3
    W = rng.standard_normal((K, D, D))
    Q = W @ W.transpose(0, 2, 1) + 0.1 * np.eye(D)
    ll = O.ar_log_likelihood(x, Ab, Q)
    assert ll.shape == (N, T - L, K)
    for n, t, k in [(0, 0, 0), (1, 3, 2), (1, 4, 3)]:
        tt = t + L
        xin = np.concatenate([x[n, tt - 2], x[n, tt - 1], [1.0]])   # oldest lag first
        r = x[n, tt] - Ab[k] @ xin
        want = -0.5 * r @ np.linalg.solve(Q[k], r) - 0.5 * np.log(np.linalg.det(2 * np.pi * Q[k]))
        assert ll[n, t, k] == pytest.approx(want, rel=1e-11)


# ------------------------------------------------------------------ Kalman -------

def _dense_lds(rng, n=4, p=3, Tp=5, K=2, masked_tail=0):
    A = rng.standard_normal((K, n, n)) * 0.4
    B = rng.standard_normal((K, n)) * 0.2
    W = rng.standard_normal((K, n, n))
    Q = W @ W.transpose(0, 2, 1) * 0.1 + 0.05 * np.eye(n)
    C = rng.standard_normal((p, n))
    D = rng.standard_normal(p)
    Rs = rng.uniform(0.3, 2.0, size=(Tp, p))
    ys = rng.standard_normal((Tp, p))
    zs = rng.integers(0, K, size=Tp - 1)
    m0 = np.zeros(n)
    S0 = np.eye(n) * 10
    mask = np.ones(Tp - 1)
    if masked_tail:
        mask[-masked_tail:] = 0
    return ys, mask, zs, m0, S0, A, B, Q, C, D, Rs


def _dense_posterior(ys, zs, m0, S0, A, B, Q, C, D, Rs):
    """Joint Gaussian over stacked x_0..x_{T'-1}, conditioned on all ys."""
    Tp, p = ys.shape
    n = m0.shape[0]
    mu = np.zeros((Tp, n))
    F = np.zeros((Tp * n, Tp * n))      # x = mu + F eps, eps ~ N(0, I) blocks
    mu[0] = m0
    F[:n, :n] = np.linalg.cholesky(S0)
    for t in range(Tp - 1):
        z = zs[t]
        mu[t + 1] = A[z] @ mu[t] + B[z]
        F[(t + 1) * n:(t + 2) * n] = A[z] @ F[t * n:(t + 1) * n]
        F[(t + 1) * n:(t + 2) * n, (t + 1) * n:(t + 2) * n] = np.linalg.cholesky(Q[z])
    Sx = F @ F.T
    H = np.kron(np.eye(Tp), C)
    Rd = np.diag(Rs.reshape(-1))
    my = H @ mu.reshape(-1) + np.tile(D, Tp)
    Sy = H @ Sx @ H.T + Rd
    yv = ys.reshape(-1)
    G = Sx @ H.T @ np.linalg.inv(Sy)
    post_m = mu.reshape(-1) + G @ (yv - my)
    post_S = Sx - G @ H @ Sx
    r = yv - my
    ll = -0.5 * r @ np.linalg.solve(Sy, r) - 0.5 * np.linalg.slogdet(2 * np.pi * Sy)[1]
    return post_m.reshape(Tp, n), post_S, ll


def test_kalman_filter_matches_dense_joint_gaussian():
    rng = np.random.default_rng(4)
    args = _dense_lds(rng)
    ys, mask, zs, m0, S0, A, B, Q, C, D, Rs = args
    ms, Ss, ll = O.kalman_filter(*args)
    pm, pS, pll = _dense_posterior(ys, zs, m0, S0, A, B, Q, C, D, Rs)
    n = m0.shape[0]
    assert np.allclose(ms[-1], pm[-1], rtol=1e-9, atol=1e-10)
    assert np.allclose(Ss[-1], pS[-n:, -n:], rtol=1e-9, atol=1e-10)
    assert ll == pytest.approx(pll, rel=1e-10)


def test_kalman_sampler_is_exact_affine_map_of_normals():
    """xs = mean + M [w; w_last]: mean and M M^T must be the dense posterior."""
    rng = np.random.default_rng(5)
    args = _dense_lds(rng)
    ys, mask, zs, m0, S0, A, B, Q, C, D, Rs = args
    Tp, n = ys.shape[0], m0.shape[0]
    zero = np.zeros((Tp - 1, n))
    base, _ = O.kalman_sample(*args, zero, np.zeros(n))
    cols = []
    for j in range(Tp * n):
        e = np.zeros(Tp * n)
        e[j] = 1
        xs, _ = O.kalman_sample(*args, e[:(Tp - 1) * n].reshape(Tp - 1, n), e[(Tp - 1) * n:])
        cols.append((xs - base).reshape(-1))
    M = np.stack(cols, axis=1)
    pm, pS, _ = _dense_posterior(ys, zs, m0, S0, A, B, Q, C, D, Rs)
    assert np.allclose(base, pm, rtol=1e-8, atol=1e-9)
    assert np.allclose(M @ M.T, pS, rtol=1e-7, atol=1e-9)


def test_kalman_masked_tail_quirk():
    """Masked steps: carry unchanged, zeros emitted, and the final update is
    unconditional (kalman.py:109-117,142-143)."""
    rng = np.random.default_rng(6)
    args = _dense_lds(rng, Tp=7, masked_tail=3)
    ys, mask, zs, m0, S0, A, B, Q, C, D, Rs = args
    n = m0.shape[0]
    w = rng.standard_normal((6, n))
    wl = rng.standard_normal(n)
    xs, _ = O.kalman_sample(*args, w, wl)
    assert np.all(xs[3:6] == 0)
    assert np.any(xs[6] != 0) and np.all(np.isfinite(xs))
    ms, Ss, _ = O.kalman_filter(*args)
    assert np.array_equal(ms[3], ms[4]) and np.array_equal(Ss[4], Ss[5])


# ------------------------------------------------------------------ statistics ---

def test_ar_sufficient_stats_match_naive_loops():
    rng = np.random.default_rng(7)
    N, T, D, L, K = 3, 20, 3, 2, 4
    x = rng.standard_normal((N, T, D))
    z = rng.integers(0, K, size=(N, T - L))
    mask = np.ones((N, T))
    mask[1, 15:] = 0
    F = O.ar_sufficient_stats(nlags=L, num_states=K, mask=mask, x=x, z=z)
    P = D * L + 1 + D
    want = np.zeros((K, P, P))
    for n in range(N):
        for t in range(L, T):
            if mask[n, t] > 0:
                f = np.concatenate([x[n, t - 2], x[n, t - 1], [1.0], x[n, t]])
                want[z[n, t - L]] += np.outer(f, f)
    assert np.allclose(F, want, rtol=1e-12, atol=1e-12)


def test_resample_regression_matches_reference_formulae():
    rng = np.random.default_rng(8)
    N, T, D, L, K = 2, 60, 2, 2, 2
    x = rng.standard_normal((N, T, D))
    z = rng.integers(0, K, size=(N, T - L))
    mask = np.ones((N, T))
    nu_0, S_0 = D + 2, 0.01 * np.eye(D)
    M_0 = np.pad(np.eye(D), ((0, 0), ((L - 1) * D, 1)))
    K_0 = 10 * np.eye(D * L + 1)
    chi2 = np.stack([rng.chisquare(nu_0 + 50 - np.arange(D)) for _ in range(K)])
    normals = rng.standard_normal((K, D * (D - 1) // 2))
    G = rng.standard_normal((K, D, D * L + 1))
    Ab, Q = O.resample_ar_params(nlags=L, num_states=K, mask=mask, x=x, z=z, nu_0=nu_0, S_0=S_0,
                                 M_0=M_0, K_0=K_0, chi2=chi2, normals=normals, G=G)
    # direct transcription of gibbs.py:127-136 for state 1 with dense one-hot masks
    x_in = O.pad_affine(O.get_lags(x, L)).reshape(-1, L * D + 1)
    x_out = x[:, L:].reshape(-1, D)
    m = (z.reshape(-1) == 1).astype(float)
    Soo = (x_out[:, :, None] * x_out[:, None, :] * m[:, None, None]).sum(0)
    Soi = (x_out[:, :, None] * x_in[:, None, :] * m[:, None, None]).sum(0)
    Sii = (x_in[:, :, None] * x_in[:, None, :] * m[:, None, None]).sum(0)
    K0i = np.linalg.inv(K_0)
    Kn = np.linalg.inv(K0i + Sii)
    Mn = (M_0 @ K0i + Soi) @ Kn
    Sn = S_0 + Soo + (M_0 @ K0i @ M_0.T - Mn @ np.linalg.inv(Kn) @ Mn.T)
    sig = O.sample_invwishart(Sn, nu_0 + m.sum(), chi2[1], normals[1])
    A = Mn + np.linalg.cholesky(sig) @ G[1] @ np.linalg.cholesky(Kn).T
    assert np.allclose(Q[1], sig, rtol=1e-10)
    assert np.allclose(Ab[1], A, rtol=1e-10, atol=1e-12)


def test_invwishart_identity():
    """T T^T == chol(S) (x^T x)^-1 chol(S)^T."""
    rng = np.random.default_rng(9)
    n, nu = 4, 9.0
    W = rng.standard_normal((n, n))
    S = W @ W.T + np.eye(n)
    chi2 = rng.chisquare(nu - np.arange(n))
    normals = rng.standard_normal(n * (n - 1) // 2)
    got = O.sample_invwishart(S, nu, chi2, normals)
    xm = np.diag(np.sqrt(chi2))
    xm[np.triu_indices(n, 1)] = normals
    ch = np.linalg.cholesky(S)
    want = ch @ np.linalg.inv(xm.T @ xm) @ ch.T
    assert np.allclose(got, want, rtol=1e-10)


def test_count_transitions_and_crp_naive():
    rng = np.random.default_rng(10)
    N, T, K = 3, 40, 5
    z = rng.integers(0, K, size=(N, T))
    mask = np.ones((N, T))
    mask[2, 25:] = 0
    counts = O.count_transitions(K, mask, z)
    want = np.zeros((K, K))
    for i in range(N):
        for j in range(T - 1):
            if mask[i, j] > 0 and mask[i, j + 1] > 0:
                want[z[i, j], z[i, j + 1]] += 1
    assert np.array_equal(counts, want)
    betas = rng.dirichlet(np.ones(K))
    randseq = rng.random(int(counts.sum()))
    m = O.sample_crp_tablecounts(3.0, counts.astype(int), betas, randseq)
    flat = counts.astype(int).ravel()
    starts = np.concatenate([[0], np.cumsum(flat)[:-1]]).reshape(K, K)
    wm = np.zeros((K, K), dtype=int)
    for i in range(K):
        for j in range(K):
            for k in range(int(counts[i, j])):
                wm[i, j] += randseq[starts[i, j] + k] < (3.0 * betas[j]) / (k + 3.0 * betas[j])
    assert np.array_equal(m, wm)


def test_get_lags_and_ar_to_lds_conventions():
    x = np.arange(2 * 6 * 2, dtype=float).reshape(2, 6, 2)
    lg = O.get_lags(x, 3)
    assert lg.shape == (2, 3, 6)
    assert np.array_equal(lg[0, 0], np.concatenate([x[0, 0], x[0, 1], x[0, 2]]))
    As = np.arange(2 * 2 * 6, dtype=float).reshape(2, 2, 6)
    A_, b_, Q_, C_ = O.ar_to_lds(As, np.ones((2, 2)), np.tile(np.eye(2), (2, 1, 1)), np.ones((3, 2)))
    assert np.array_equal(A_[0, :4, 2:], np.eye(4)) and np.array_equal(A_[1, -2:], As[1])
    assert np.allclose(np.diag(Q_[0]), [1e-2] * 4 + [1, 1])
    assert np.array_equal(b_[0], [0, 0, 0, 0, 1, 1]) and np.array_equal(C_[:, :4], np.zeros((3, 4)))


def test_vector_to_angle_is_atan2():
    rng = np.random.default_rng(11)
    V = rng.standard_normal((100, 2))
    assert np.allclose(O.vector_to_angle(V), np.arctan2(V[:, 1], V[:, 0]), atol=1e-8)


def test_vonmises_contract_distribution():
    rng = np.random.default_rng(12)
    S, R = 40_000, 12
    kappa = 4.0
    h = O.sample_vonmises(np.zeros(S), np.full(S, kappa), rng.random((S, R, 3)))
    from scipy.special import i0, i1
    assert abs(np.mean(np.cos(h)) - i1(kappa) / i0(kappa)) < 5e-3
    assert abs(np.mean(np.sin(h))) < 1e-2
