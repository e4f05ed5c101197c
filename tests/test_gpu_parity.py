"""GPU parity tests: every CUDA entry point, called through the C-ABI (via the ctypes
binding the product uses), against the float64 oracle on identical seeded inputs and
identical injected draws.

Tolerances (BASELINE.json north_star): marginal log-likelihoods 1e-5 relative; state
sequences >= 99.9 % agreement; sufficient statistics 1e-5 relative; integer work exact.
Where the CUDA path computes in float64 the tests are far tighter than that.
"""
import itertools

import numpy as np
import pytest

import kpms_helpers as H
from oracle import kpms_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    import torch
    assert torch.cuda.is_available()
    from keypointmoseq_b200 import gibbs
    return gibbs


def T_(a, dtype=None):
    import torch
    t = torch.as_tensor(np.ascontiguousarray(a), device="cuda")
    return t if dtype is None else t.to(dtype)


def N_(t):
    return t.detach().cpu().numpy()


# ------------------------------------------------------------------ RNG fills
def test_fill_kernels_moments():
    import torch
    from keypointmoseq_b200 import distributions as dist
    u = dist.fill_uniform((1 << 20,), 123, 1, "cuda")
    assert float(u.min()) >= 0.0 and float(u.max()) < 1.0
    assert abs(float(u.mean()) - 0.5) < 2e-3 and abs(float(u.var()) - 1 / 12) < 2e-3
    w = dist.fill_normal((1 << 20,), 123, 2, "cuda")
    assert abs(float(w.mean())) < 5e-3 and abs(float(w.var()) - 1.0) < 1e-2
    assert abs(float((w ** 4).mean()) - 3.0) < 0.1
    g = dist.fill_gamma((1 << 20,), 4.0, 123, 3, "cuda")
    assert abs(float(g.mean()) - 4.0) < 2e-2 and abs(float(g.var()) - 4.0) < 1e-1
    assert float(g.min()) > 0
    u2 = dist.fill_uniform((1 << 20,), 123, 1, "cuda")
    assert torch.equal(u, u2)                      # deterministic in (seed, stream)
    u3 = dist.fill_uniform((1001,), 7, 1, "cuda")  # ragged tail path
    assert u3.shape == (1001,) and float(u3.max()) < 1.0


# ------------------------------------------------------------------ a1
@pytest.mark.parametrize("D,L,K", [(10, 3, 7), (7, 3, 5), (2, 2, 3), (10, 3, 100)])
@pytest.mark.parametrize("impl", [1, 0])
def test_ar_loglik_parity(G, D, L, K, impl):
    d, s, p, Gm = H.problem(seed=1, N=3, T=70, D=D, L=L, K=K, ragged=False)
    want = O.ar_log_likelihood(s['x'], p['Ab'], p['Q'])
    got = N_(G.ar_log_likelihood(x=T_(s['x']), Ab=T_(p['Ab']), Q=T_(p['Q']), impl=impl)).astype(np.float64)
    assert got.shape == want.shape
    err = np.abs(got - want) / np.maximum(1.0, np.abs(want))
    assert err.max() < 2e-5, err.max()


# ------------------------------------------------------------------ a2 + a3
def _hmm_case(seed, N, T, K, ragged=True, sticky=True):
    rng = np.random.default_rng(seed)
    pi = rng.dirichlet(np.ones(K), size=K)
    if sticky:
        pi = 0.1 * pi + 0.9 * np.eye(K)
    ll = H.f32r(rng.standard_normal((N, T, K)) * 3 - 4)
    mask = np.ones((N, T))
    if ragged:
        for n in range(1, N):
            mask[n, rng.integers(T // 2, T):] = 0
    u = H.f32r(rng.random((N, T)))
    return ll, mask, pi, u


@pytest.mark.parametrize("K", [3, 33, 50, 96, 100, 128, 200])
def test_hmm_ffbs_parity(G, K):
    ll, mask, pi, u = _hmm_case(K, N=4, T=400, K=K)
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    z, lg = G.sample_hmm_stateseqs(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u))
    z, lg = N_(z), N_(lg)
    assert np.all(np.abs(lg - lo) <= 1e-5 * np.abs(lo)), (lg, lo)
    valid = mask > 0
    agree = (z[valid] == zo[valid]).mean()
    assert agree >= 0.999, agree
    assert np.all(z[~valid] == 0)


@pytest.mark.parametrize("K", [3, 33, 50, 98, 100])
def test_hmm_chunked_tensor_core_any_state_count(G, K):
    """Chunked tcgen05 filter for state counts that are not multiples of 4 (states are padded with
    zero-probability ones inside the library), e.g. K = 50 of BASELINE configs[2]."""
    ll, mask, pi, u = _hmm_case(100 + K, N=4, T=2400, K=K)
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    z, lg, info = G.sample_hmm_stateseqs(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), chunk_len=300, burn_in=256,
                                         strict=False, return_info=True)
    assert float(info["boundary_err"]) < 1e-4
    z, lg = N_(z), N_(lg)
    assert np.all(np.abs(lg - lo) <= 1e-5 * np.abs(lo)), (lg, lo)
    valid = mask > 0
    assert (z[valid] == zo[valid]).mean() >= 0.999 and z.max() < K
    assert np.all(z[~valid] == 0)


@pytest.mark.parametrize("K", [130, 200, 256, 500])
def test_hmm_chunked_large_state_counts(G, options, K):
    """128 < K <= 512: chunk-parallel filter with 32 chains per CTA and transition tiles streamed from
    L2 (hmm_big.cu), sampler without stay bits; against the oracle and against the sequential
    K-generic kernel."""
    import torch
    ll, mask, pi, u = _hmm_case(300 + K, N=3, T=1500, K=K)
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    kw = dict(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), return_info=True)
    z, lg, info = G.sample_hmm_stateseqs(chunk_len=300, burn_in=256, strict=False, **kw)
    assert float(info["boundary_err"]) < 1e-4
    z_seq, lg_seq, _ = G.sample_hmm_stateseqs(chunk_len=0, **kw)
    valid = mask > 0
    assert np.all(np.abs(N_(lg) - lo) <= 1e-5 * np.abs(lo)), (N_(lg), lo)
    assert (N_(z)[valid] == zo[valid]).mean() >= 0.999 and int(z.max()) < K
    assert np.all(N_(z)[~valid] == 0)
    assert float((z != z_seq)[torch.as_tensor(valid, device="cuda")].float().mean()) < 1e-3
    # a warm-up that is too short is repaired on the device (chunks re-filtered from their predecessor's end;
    # five chunks per session: a cascade of four rounds) ...
    z_r, lg_r, rep = G.sample_hmm_stateseqs(chunk_len=300, burn_in=1, strict=False, **kw)
    assert float(rep["boundary_err"]) < 1e-4
    assert np.all(np.abs(N_(lg_r) - lo) <= 1e-5 * np.abs(lo))
    assert float((z_r != z_seq)[torch.as_tensor(valid, device="cuda")].float().mean()) < 1e-3
    # ... and shows up in the reported mismatch when the repair is switched off
    options("hmm_repair_rounds", 0)
    _, _, bad = G.sample_hmm_stateseqs(chunk_len=300, burn_in=1, strict=False, **kw)
    assert float(bad["boundary_err"]) > 1e-3


def test_hmm_marginal_loglik_known_answer(G):
    """Brute-force enumeration of all K^T paths (K=3, T=6), through the CUDA kernel."""
    rng = np.random.default_rng(0)
    K, T = 3, 6
    pi = rng.dirichlet(np.ones(K), size=K)
    ll = H.f32r(rng.standard_normal((1, T, K)) * 3 - 5)
    logp = []
    for zp in itertools.product(range(K), repeat=T):
        lp = -np.log(K) + ll[0, 0, zp[0]]
        for t in range(1, T):
            lp += np.log(pi[zp[t - 1], zp[t]]) + ll[0, t, zp[t]]
        logp.append(lp)
    want = np.logaddexp.reduce(np.array(logp))
    _, lg = G.sample_hmm_stateseqs(ll=T_(ll), mask=T_(np.ones((1, T))), pi=T_(pi), u=T_(np.full((1, T), .5)))
    assert abs(float(lg[0]) - want) <= 1e-5 * abs(want)


def test_hmm_all_masked_and_single_frame(G):
    K = 5
    pi = np.full((K, K), 1.0 / K)
    z, lg = G.sample_hmm_stateseqs(ll=T_(np.zeros((2, 9, K))), mask=T_(np.zeros((2, 9))), pi=T_(pi),
                                   u=T_(np.full((2, 9), .3)))
    assert np.all(N_(z) == 0) and np.all(N_(lg) == 0.0)
    z, lg = G.sample_hmm_stateseqs(ll=T_(np.zeros((1, 1, K))), mask=T_(np.ones((1, 1))), pi=T_(pi),
                                   u=T_(np.full((1, 1), .55)))
    assert int(z[0, 0]) == 2 and abs(float(lg[0])) < 1e-6   # uniform posterior, u=.55 -> state 2


@pytest.mark.parametrize("D,L,K", [(10, 3, 20), (7, 3, 100)])
def test_resample_stateseqs_parity(G, D, L, K):
    d, s, p, Gm = H.problem(seed=3, N=4, T=300, D=D, L=L, K=K, ragged=True)
    rng = np.random.default_rng(5)
    u = H.f32r(rng.random((4, 300 - L)))
    zo, lo = O.resample_stateseqs(x=s['x'], mask=d['mask'], Ab=p['Ab'], Q=p['Q'], pi=p['pi'], u=u)
    z, lg = G.resample_stateseqs(0, x=T_(s['x']), mask=T_(d['mask']), Ab=T_(p['Ab']), Q=T_(p['Q']),
                                 pi=T_(p['pi']), draws={'u': T_(u)})
    z, lg = N_(z), N_(lg)
    assert np.all(np.abs(lg - lo) <= 1e-5 * np.abs(lo)), (lg, lo)
    valid = d['mask'][:, L:] > 0
    assert (z[valid] == zo[valid]).mean() >= 0.999


# ------------------------------------------------------------------ a5-a7
def _latents_case(seed, N, T, k, dk, D, L, K, ragged=True):
    d, s, p, Gm = H.problem(seed=seed, N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, ragged=ragged)
    rng = np.random.default_rng(seed + 100)
    dr = H.sweep_draws(rng, N, T, k, dk, D, L, K)
    return d, s, p, Gm, dr


@pytest.mark.parametrize("k,D,L,K", [(8, 10, 3, 6), (16, 10, 3, 4), (8, 7, 3, 5), (5, 2, 2, 3), (6, 10, 2, 3)])
def test_resample_latents_parity(G, k, D, L, K):
    N, T, dk = 3, 90, 2
    d, s, p, Gm, dr = _latents_case(11, N, T, k, dk, D, L, K)
    xo, llo = O.resample_latents(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], z=s['z'], s=s['s'], Cd=p['Cd'],
                                 sigmasq=p['sigmasq'], Ab=p['Ab'], Q=p['Q'], Gamma=Gm, w=dr['x_w'],
                                 w_last=dr['x_w_last'], return_loglik=True)
    x, info = G.resample_latents(0, Y=T_(d['Y']), mask=T_(d['mask']), v=T_(s['v']), h=T_(s['h']),
                                 z=T_(s['z']), s=T_(s['s']), Cd=T_(p['Cd']), sigmasq=T_(p['sigmasq']),
                                 Ab=T_(p['Ab']), Q=T_(p['Q']), Gamma=T_(Gm),
                                 draws={'w': T_(dr['x_w']), 'w_last': T_(dr['x_w_last'])}, return_info=True)
    x, ll = N_(x).astype(np.float64), N_(info['loglik'])
    assert np.all(np.isfinite(x))
    assert np.abs(x - xo).max() < 2e-6 * max(1.0, np.abs(xo).max()), np.abs(x - xo).max()
    assert np.all(np.abs(ll - llo) <= 1e-9 * np.abs(llo)), (ll, llo)


def test_resample_latents_chunked_matches_exact(G):
    N, T, k, dk, D, L, K = 3, 700, 8, 2, 10, 3, 5
    d, s, p, Gm, dr = _latents_case(12, N, T, k, dk, D, L, K)
    kw = dict(Y=T_(d['Y']), mask=T_(d['mask']), v=T_(s['v']), h=T_(s['h']), z=T_(s['z']), s=T_(s['s']),
              Cd=T_(p['Cd']), sigmasq=T_(p['sigmasq']), Ab=T_(p['Ab']), Q=T_(p['Q']), Gamma=T_(Gm),
              draws={'w': T_(dr['x_w']), 'w_last': T_(dr['x_w_last'])}, return_info=True)
    x0, i0 = G.resample_latents(0, chunk_len=0, **kw)
    x1, i1 = G.resample_latents(0, chunk_len=128, burn_in=160, strict=False, **kw)
    berr = N_(i1['boundary_err'])
    assert berr.max() < 1e-9, berr
    assert float((x0 - x1).abs().max()) < 1e-6
    l0, l1 = N_(i0['loglik']), N_(i1['loglik'])
    assert np.all(np.abs(l0 - l1) <= 1e-9 * np.abs(l0))
    # too short a warm-up must be visible in the reported mismatch
    _, i2 = G.resample_latents(0, chunk_len=128, burn_in=2, strict=False, **kw)
    assert N_(i2['boundary_err']).max() > 1e-6
    # strict mode notices the mismatch and redoes the call (longer warm-up, then unchunked) until the
    # mismatch is at rounding level
    x3, i3 = G.resample_latents(0, chunk_len=128, burn_in=2, strict=True, **kw)
    assert float((x0 - x3).abs().max()) < 1e-6 and N_(i3['boundary_err']).max() < 1e-12


# ------------------------------------------------------------------ a8-a11
@pytest.mark.parametrize("D,L,K,T", [(10, 3, 6, 200), (7, 3, 100, 500), (2, 2, 3, 50)])
def test_ar_stats_parity(G, D, L, K, T):
    d, s, p, Gm = H.problem(seed=21, N=4, T=T, D=D, L=L, K=K, ragged=True)
    want = O.ar_sufficient_stats(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=s['z'])
    got = N_(G.ar_sufficient_stats(nlags=L, num_states=K, mask=T_(d['mask']), x=T_(s['x']), z=T_(s['z'])))
    scale = np.abs(want).max(axis=(1, 2), keepdims=True) + 1e-300
    assert (np.abs(got - want) / scale).max() < 1e-12
    assert np.array_equal(got, got.transpose(0, 2, 1))


def test_resample_ar_params_injected_draws(G):
    D, L, K = 10, 3, 5
    d, s, p, Gm = H.problem(seed=22, N=3, T=400, D=D, L=L, K=K, ragged=True)
    hp = H.hyp(D, L, K)
    rng = np.random.default_rng(1)
    F = O.ar_sufficient_stats(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=s['z'])
    cnt = F[:, D * L, D * L]
    chi2 = np.stack([rng.chisquare(hp['nu_0'] + cnt[i] - np.arange(D)) for i in range(K)])
    normals = rng.standard_normal((K, D * (D - 1) // 2))
    Gd = rng.standard_normal((K, D, D * L + 1))
    Abo, Qo = O.resample_ar_params(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=s['z'], nu_0=hp['nu_0'],
                                   S_0=hp['S_0'], M_0=hp['M_0'], K_0=hp['K_0'], chi2=chi2, normals=normals, G=Gd)
    Ab, Q = G.resample_ar_params(0, nlags=L, num_states=K, mask=T_(d['mask']), x=T_(s['x']), z=T_(s['z']),
                                 nu_0=hp['nu_0'], S_0=T_(hp['S_0']), M_0=T_(hp['M_0']), K_0=T_(hp['K_0']),
                                 draws={'chi2': T_(chi2), 'normals': T_(normals), 'G': T_(Gd)})
    assert np.allclose(N_(Q), Qo, rtol=1e-7, atol=1e-12)
    assert np.allclose(N_(Ab), Abo, rtol=1e-6, atol=1e-9)


@pytest.mark.parametrize("K", [5, 100, 210])
def test_transition_counts_exact(G, K):
    rng = np.random.default_rng(K)
    N, T, L = 5, 300, 3
    z = rng.integers(0, K, size=(N, T - L))
    mask = np.ones((N, T))
    mask[1, 200:] = 0
    mask[2, 5:] = 0
    want = O.count_transitions(K, mask[:, -z.shape[1]:], z)
    got = N_(G.transition_counts(z=T_(z), mask=T_(mask), num_states=K))
    assert got.dtype == np.int64 and np.array_equal(got, want.astype(np.int64))


def test_crp_tablecounts_exact(G):
    import torch
    rng = np.random.default_rng(31)
    K = 12
    counts = rng.integers(0, 40, size=(K, K))
    counts[np.arange(K), np.arange(K)] += rng.integers(200, 3000, size=K)
    counts[3] = 0
    betas = rng.dirichlet(np.ones(K))
    alpha, kappa = 5.7, 40.0
    randseq = H.f32r(rng.random(int(counts.sum())))
    u_thin = H.f32r(rng.random((K, 4096)))
    want_raw = O.sample_crp_tablecounts(alpha, counts, betas, randseq)
    # default = what the reference returns (its thinning is lost in a temporary): the raw counts
    ms0, ms_raw0 = G.sample_ms(0, T_(counts), T_(betas), alpha, kappa, draws={'randseq': T_(randseq)})
    assert np.array_equal(N_(ms0), want_raw) and np.array_equal(N_(ms_raw0), want_raw)
    assert np.array_equal(O.sample_ms(counts, betas, alpha, kappa, randseq), want_raw)
    # thin=True: the intended binomial thinning of the sticky diagonal, as Bernoulli sums
    want = O.sample_ms(counts, betas, alpha, kappa, randseq, u_thin, thin=True)
    ms, ms_raw = G.sample_ms(0, T_(counts), T_(betas), alpha, kappa,
                             draws={'randseq': T_(randseq), 'u_thin': T_(u_thin)}, thin=True)
    assert np.array_equal(N_(ms_raw), want_raw) and np.array_equal(N_(ms), want)
    assert (want.diagonal() < want_raw.diagonal()).any()
    # injected draws that are too short are refused loudly, not silently capped
    from keypointmoseq_b200._lib import KpmsError
    with pytest.raises(KpmsError):
        G.sample_ms(0, T_(counts), T_(betas), alpha, kappa, draws={'randseq': T_(randseq[:-5]), 'u_thin': T_(u_thin)})
    with pytest.raises(KpmsError):
        G.sample_ms(0, T_(counts), T_(betas), alpha, kappa, draws={'randseq': T_(randseq), 'u_thin': T_(u_thin[:, :2])},
                    thin=True)
    # Dirichlet concentrations of transitions.py:38-39 from the same statistics
    gb = rng.gamma(want.sum(0) + 1e3 / K)
    betas_new = gb / gb.sum()
    conc = alpha * betas_new[None] + counts + kappa * np.eye(K)
    gp = rng.gamma(conc)
    b, pi = G.sample_hdp_transitions(0, T_(counts), T_(betas), alpha, kappa, 1e3, thin=True,
                                     draws={'randseq': T_(randseq), 'u_thin': T_(u_thin), 'betas_g': T_(gb),
                                            'pi_g': T_(gp)})
    assert np.allclose(N_(b), betas_new, rtol=1e-13) and np.allclose(N_(pi), gp / gp.sum(1, keepdims=True), rtol=1e-13)


def test_crp_philox_equals_filled_uniforms(G):
    """The in-kernel Philox draws are the numbers kpms_fill_uniform would have written: the table
    counts from kpms_crp_tablecounts_philox equal those from pre-filled uniforms, bit for bit
    (raw counts and, with thin=True, the thinned diagonal)."""
    import torch
    from keypointmoseq_b200 import distributions as dist
    rng = np.random.default_rng(32)
    K = 9
    counts = rng.integers(0, 50, size=(K, K))
    counts[np.arange(K), np.arange(K)] += rng.integers(500, 20000, size=K)
    counts[2] = 0
    betas = rng.dirichlet(np.ones(K))
    alpha, kappa, seed = 100.0, 300.0, 77
    randseq = dist.fill_uniform((int(counts.sum()),), seed, dist.STREAM_CRP, "cuda")
    cap = 4096
    u_thin = torch.empty((K, cap), device="cuda")
    for i in range(K):   # row i of the thinning draws = numbers (i << 32) .. of the stream
        u_thin[i] = dist.fill_uniform((1, cap), seed, dist.STREAM_THIN, "cuda", first_row=(i << 32) // cap)[0]
    a, a_raw = G.sample_ms(seed, T_(counts), T_(betas), alpha, kappa, thin=True)
    b, b_raw = G.sample_ms(seed, T_(counts), T_(betas), alpha, kappa, thin=True, draws={'randseq': randseq, 'u_thin': u_thin})
    assert torch.equal(a_raw, b_raw) and torch.equal(a, b)
    assert int(a_raw.sum()) > 0 and (N_(a).diagonal() < N_(a_raw).diagonal()).any()
    c, c_raw = G.sample_ms(seed, T_(counts), T_(betas), alpha, kappa)      # reference behaviour: no thinning
    assert torch.equal(c, a_raw) and torch.equal(c_raw, a_raw)


def test_resample_hdp_transitions_philox_is_valid(G):
    d, s, p, Gm = H.problem(seed=33, N=3, T=500, K=8, ragged=True)
    betas, pi = G.resample_hdp_transitions(7, z=T_(s['z']), mask=T_(d['mask']), betas=T_(p['betas']),
                                           alpha=100.0, kappa=2e4, gamma=1e3, num_states=8)
    betas, pi = N_(betas), N_(pi)
    assert np.all(betas > 0) and abs(betas.sum() - 1) < 1e-12
    assert np.all(pi >= 0) and np.allclose(pi.sum(1), 1.0) and np.all(np.diag(pi) > 0.9)


# ------------------------------------------------------------------ a13-a16
@pytest.mark.parametrize("k,dk,D", [(8, 2, 10), (16, 2, 10), (5, 3, 7)])
def test_obs_passes_parity(G, k, dk, D):
    N, T, L, K = 3, 120, 3, 4
    d, s, p, Gm = H.problem(seed=41, N=N, T=T, k=k, dk=2, D=D, L=L, K=K, ragged=True)
    rng = np.random.default_rng(2)
    if dk == 3:   # 3-D keypoints: synthesise a third coordinate
        d['Y'] = H.f32r(np.concatenate([d['Y'], rng.standard_normal((N, T, k, 1))], -1))
        s['v'] = H.f32r(np.concatenate([s['v'], rng.standard_normal((N, T, 1))], -1))
        p['Cd'] = rng.standard_normal(((k - 1) * 3, D + 1)) / 3
    sig = rng.uniform(0.5, 2.0, size=k)
    common = dict(Y=T_(d['Y']), mask=T_(d['mask']), v=T_(s['v']), h=T_(s['h']), x=T_(s['x']), s=T_(s['s']),
                  Cd=T_(p['Cd']), sigmasq=T_(sig), Gamma=T_(Gm))
    oc = dict(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], x=s['x'], s=s['s'], Cd=p['Cd'], sigmasq=sig, Gamma=Gm)
    # a13
    So, mo = O.obs_variance_stats(**oc)
    S, m = G.obs_variance_stats(**common)
    assert np.allclose(N_(S), So, rtol=1e-11) and float(m) == mo
    hp = H.hyp(D, L, K)
    g = rng.gamma((hp['nu_sigma'] + 3 * mo) / 2, size=k)
    want = O.resample_obs_variance(**oc, nu_sigma=hp['nu_sigma'], sigmasq_0=hp['sigmasq_0'], g=g)
    got = N_(G.resample_obs_variance(0, **common, nu_sigma=hp['nu_sigma'], sigmasq_0=hp['sigmasq_0'],
                                     draws={'g': T_(g)}))
    assert np.allclose(got, want, rtol=1e-11)
    # a14
    g = H.f32r(rng.gamma((hp['nu_s'] + 3) / 2, size=(N, T, k)))
    want = O.resample_scales(**oc, nu_s=hp['nu_s'], s_0=hp['s_0'], g=g)
    got = N_(G.resample_scales(0, **common, nu_s=hp['nu_s'], s_0=hp['s_0'], draws={'g': T_(g)})).astype(np.float64)
    assert np.allclose(got, want, rtol=3e-7)
    # a15
    u = H.f32r(rng.random((N, T, 10, 3)) * 0.998 + 0.001)
    tho, kao = O.heading_stats(**oc)
    ho = O.sample_vonmises(tho, kao, u)
    hh, info = G.resample_heading(0, **common, draws={'u': T_(u)}, return_info=True)
    ka = N_(info['kappa'])
    assert np.allclose(ka, kao, rtol=1e-10)
    dth = np.angle(np.exp(1j * (N_(info['theta']) - tho)))
    assert np.abs(dth).max() < 1e-9
    dh = np.angle(np.exp(1j * (N_(hh).astype(np.float64) - ho)))
    assert np.abs(dh).max() < 1e-6, np.abs(dh).max()
    # a16
    w = H.f32r(rng.standard_normal((N, T - 1, dk)))
    wl = H.f32r(rng.standard_normal((N, dk)))
    vo, llo = O.resample_location(**oc, sigmasq_loc=0.5, w=w, w_last=wl, return_loglik=True)
    vv, info = G.resample_location(0, **common, sigmasq_loc=0.5, draws={'w': T_(w), 'w_last': T_(wl)},
                                   return_info=True)
    vv = N_(vv).astype(np.float64)
    assert np.abs(vv - vo).max() < 3e-7 * max(1.0, np.abs(vo).max())
    assert np.all(np.abs(N_(info['loglik']) - llo) <= 1e-9 * np.abs(llo))


# ------------------------------------------------------------------ whole sweep
def test_full_sweep_runs_with_library_rng(G):
    import torch
    N, T, k, dk, D, L, K = 4, 400, 8, 2, 10, 3, 10
    d, s, p, Gm = H.problem(seed=51, N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, ragged=True)
    hp = H.hyp(D, L, K)
    data = {a: T_(b, torch.float32) for a, b in d.items()}
    st = {a: (T_(b, torch.float32) if a != 'z' else T_(b, torch.int32)) for a, b in s.items()}
    pr = {a: T_(b) for a, b in p.items()}
    hpt = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in hp.items()}
    for it in range(2):
        key = (0, it)
        pr['Ab'], pr['Q'] = G.resample_ar_params(key, **data, **st, **pr, **hpt)
        pr['sigmasq'] = G.resample_obs_variance(key, **data, **st, **pr, **hpt, Gamma=T_(Gm))
        pr['betas'], pr['pi'] = G.resample_hdp_transitions(key, **data, **st, **pr, **hpt)
        st['z'] = G.resample_stateseqs(key, **data, **st, **pr)[0]
        st['x'] = G.resample_latents(key, **data, **st, **pr, Gamma=T_(Gm))
        st['h'] = G.resample_heading(key, **data, **st, **pr, Gamma=T_(Gm))
        st['v'] = G.resample_location(key, **data, **st, **pr, **hpt, Gamma=T_(Gm))
        st['s'] = G.resample_scales(key, **data, **st, **pr, **hpt, Gamma=T_(Gm))
    torch.cuda.synchronize()
    for name in ('x', 'v', 'h', 's'):
        assert torch.isfinite(st[name]).all(), name
    assert st['z'].dtype == torch.int32 and int(st['z'].min()) >= 0 and int(st['z'].max()) < K
    assert torch.isfinite(pr['Ab']).all() and torch.isfinite(pr['Q']).all()
    assert float(st['s'].min()) > 0 and float(pr['sigmasq'].min()) > 0
    # latents should still track the data: residual variance stays near the noise level
    S_y, m = G.obs_variance_stats(**data, **st, **pr, Gamma=T_(Gm))
    assert float((S_y / (2 * m)).max()) < 5.0


def test_host_pipelined_sweep_equals_device_sweep(G):
    """parallel.gibbs_sweep_host (uploads ordered by first use, observation statistics deferred
    past the state sequences, downloads as states become final) must give the very same states
    and parameters as parallel.gibbs_sweep on device-resident inputs."""
    import torch
    from keypointmoseq_b200 import parallel
    N, T, k, dk, D, L, K = 3, 300, 8, 2, 10, 3, 10
    d, s, p, Gm = H.problem(seed=77, N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, ragged=True)
    hp = H.hyp(D, L, K)
    hpt = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in hp.items()}
    host_d = {a: torch.as_tensor(b, dtype=torch.float32).pin_memory() for a, b in d.items()}
    host_s = {a: (torch.as_tensor(b, dtype=torch.float32) if a != 'z' else torch.as_tensor(b.astype(np.int32))).pin_memory()
              for a, b in s.items()}
    pr = {a: T_(b) for a, b in p.items()}
    dev_d = {a: b.cuda() for a, b in host_d.items()}
    dev_s = {a: b.cuda() for a, b in host_s.items()}
    st1, pr1 = parallel.gibbs_sweep((3, 1), dev_d, dev_s, pr, hpt, Gamma=T_(Gm))
    out, pr2, st2 = parallel.gibbs_sweep_host((3, 1), host_d, host_s, pr, hpt, Gamma=T_(Gm))
    torch.cuda.synchronize()
    for name in ('z', 'x', 'h', 'v', 's'):
        assert torch.equal(st1[name].cpu(), out[name]), name
        assert torch.equal(st2[name].cpu(), out[name]), name
    for name in pr1:
        assert torch.equal(pr1[name], pr2[name]), name
    # a second call reuses the output buffers
    out2, _, _ = parallel.gibbs_sweep_host((3, 1), host_d, host_s, pr, hpt, Gamma=T_(Gm), out=out)
    assert out2 is out and torch.equal(st1['x'].cpu(), out['x'])


def test_host_sweep_session_groups(G):
    """A loop of host sweeps that feeds every sweep the previous one's output tensors runs the pose
    steps session group by session group (uploads, kernels and downloads overlapping).  Against the
    ungrouped sweep from the same inputs: state sequences and parameters identical (the carried
    observation statistics are bit-identical to recomputed ones), continuous states equal up to the
    Kalman chunk layout, which is chosen per launch."""
    import torch
    from keypointmoseq_b200 import parallel
    N, T, k, dk, D, L, K = 6, 1500, 8, 2, 10, 3, 10
    d, s, p, Gm = H.problem(seed=78, N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, ragged=True)
    hpt = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in H.hyp(D, L, K).items()}
    host_d = {a: torch.as_tensor(b, dtype=torch.float32).pin_memory() for a, b in d.items()}
    host_s = {a: (torch.as_tensor(b, dtype=torch.float32) if a != 'z' else torch.as_tensor(b.astype(np.int32))).pin_memory()
              for a, b in s.items()}
    pr = {a: T_(b) for a, b in p.items()}
    out1, pr1, _ = parallel.gibbs_sweep_host((3, 1), host_d, host_s, pr, hpt, Gamma=T_(Gm), groups=3)
    assert "_carry" in out1                       # first sweep of the loop: ungrouped, leaves its statistics
    # second sweep, twice from the same inputs: grouped (out1 passed back untouched) and ungrouped
    outg, prg, stg = parallel.gibbs_sweep_host((3, 2), host_d, out1, pr1, hpt, Gamma=T_(Gm), groups=3)
    plain = {a: b.clone().pin_memory() for a, b in out1.items() if not a.startswith("_")}
    outp, prp, stp = parallel.gibbs_sweep_host((3, 2), host_d, plain, pr1, hpt, Gamma=T_(Gm), groups=3)
    torch.cuda.synchronize()
    assert torch.equal(outg['z'], outp['z'])
    for name in prg:
        assert torch.equal(prg[name], prp[name]), name
    for name in ('x', 'h', 'v', 's'):
        a, b = outg[name].double(), outp[name].double()
        tol = 2e-6 * max(1.0, float(b.abs().max())) if name != 's' else 1e-5 * float(b.abs().max())
        if name == 'h':                            # angles: compare on the circle
            dlt = torch.remainder(a - b + np.pi, 2 * np.pi) - np.pi
            assert float(dlt.abs().max()) < 1e-4, name
        else:
            assert float((a - b).abs().max()) < tol, (name, float((a - b).abs().max()))
        assert torch.equal(stg[name].cpu(), outg[name]), name      # every group's part reached the host buffer
    # a caller who changes a state in place gets the ungrouped path again (and correct results)
    outg['v'].mul_(1.0)
    out3, _, _ = parallel.gibbs_sweep_host((3, 3), host_d, outg, prg, hpt, Gamma=T_(Gm), groups=3)
    ref3 = {a: b.clone().pin_memory() for a, b in outg.items() if not a.startswith("_")}
    out4, _, _ = parallel.gibbs_sweep_host((3, 3), host_d, ref3, prg, hpt, Gamma=T_(Gm), groups=1)
    for name in ('z', 'x', 'h', 'v', 's'):
        assert torch.equal(out3[name], out4[name]), name


def test_cpu_tensor_is_rejected(G):
    import torch
    from keypointmoseq_b200._lib import KpmsError
    with pytest.raises(KpmsError):
        G.ar_log_likelihood(x=torch.zeros(1, 10, 10), Ab=torch.zeros(2, 10, 31), Q=torch.eye(10).repeat(2, 1, 1))


@pytest.mark.parametrize("sigmasq_loc,sig_scale,T", [(0.5, 1.0, 900), (1e-4, 50.0, 9000), (0.5, 1.0, 36000)])
def test_location_exact_time_parallel_scan(G, options, sigmasq_loc, sig_scale, T):
    """The location sampler is an exact associative scan over time (Moebius maps for the filter
    precision, affine maps for the mean and the backward samples): it equals the sequential oracle for
    any model, including the slowly forgetting one (small sigmasq_loc, large observation noise) in
    which a chunk scheme with a fixed 128-step warm-up visibly fails."""
    N, k, dk, D, L, K = 2, 8, 2, 10, 3, 4
    d, s, p, Gm = H.problem(seed=61, N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, ragged=True)
    rng = np.random.default_rng(3)
    w = H.f32r(rng.standard_normal((N, T - 1, dk)))
    wl = H.f32r(rng.standard_normal((N, dk)))
    oc = dict(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], x=s['x'], s=s['s'], Cd=p['Cd'],
              sigmasq=p['sigmasq'] * sig_scale, Gamma=Gm)
    vo, llo = O.resample_location(**oc, sigmasq_loc=sigmasq_loc, w=w, w_last=wl, return_loglik=True)
    kw = dict(**{a: T_(b) for a, b in oc.items()}, sigmasq_loc=sigmasq_loc, draws={'w': T_(w), 'w_last': T_(wl)},
              return_info=True)
    vv, info = G.resample_location(0, **kw)
    err = np.abs(N_(vv).astype(np.float64) - vo).max() / max(1.0, np.abs(vo).max())
    assert err < 3e-7, err
    assert np.all(np.abs(N_(info['loglik']) - llo) <= 1e-9 * np.abs(llo))
    # the warm-up variant kept behind an option: fine on the benign model, off on the slow one
    options("location_impl", 1)
    vw, _ = G.resample_location(0, **kw)
    errw = np.abs(N_(vw).astype(np.float64) - vo).max() / max(1.0, np.abs(vo).max())
    assert (errw < 3e-7) if sigmasq_loc > 0.1 else (errw > 1e-4), errw


def test_hmm_long_sequence_bench_shape(G):
    """K=100 at a few thousand frames: marginal log-likelihood and state agreement at scale."""
    ll, mask, pi, u = _hmm_case(7, N=3, T=6000, K=100, ragged=True)
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    z, lg = G.sample_hmm_stateseqs(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u))
    z, lg = N_(z), N_(lg)
    assert np.all(np.abs(lg - lo) <= 1e-5 * np.abs(lo))
    valid = mask > 0
    assert (z[valid] == zo[valid]).mean() >= 0.999


def test_hmm_chunked_forward_matches_exact(G):
    """Chunked forward filter (independent CTAs with warm-up) against the sequential one."""
    import torch
    ll, mask, pi, u = _hmm_case(11, N=3, T=9000, K=100, ragged=True)
    kw = dict(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), return_info=True)
    z0, l0, i0 = G.sample_hmm_stateseqs(chunk_len=0, **kw)
    z1, l1, i1 = G.sample_hmm_stateseqs(chunk_len=1500, burn_in=512, strict=False, **kw)
    assert float(i0['boundary_err']) == 0.0 and float(i1['boundary_err']) < 1e-5
    assert float((z0 != z1).float().mean()) < 1e-3
    assert np.all(np.abs(N_(l0) - N_(l1)) <= 1e-5 * np.abs(N_(l0)))   # north_star tolerance
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    valid = mask > 0
    assert (N_(z1)[valid] == zo[valid]).mean() >= 0.999
    # a uselessly short warm-up is reported, and strict mode repairs it
    z2, l2, i2 = G.sample_hmm_stateseqs(chunk_len=1500, burn_in=0, strict=False, **kw)
    assert float(i2['boundary_err']) > 1e-3
    z3, l3, i3 = G.sample_hmm_stateseqs(chunk_len=1500, burn_in=0, strict=True, **kw)
    assert float((z3 != z0).float().mean()) < 1e-3 and float(i3['boundary_err']) < 1e-5


@pytest.mark.parametrize("K", [8, 32, 50, 60, 64, 100, 104, 128])
def test_hmm_chunked_all_instantiations(G, K):
    """Chunk-parallel filter + sampler for every kernel variant: K % 4 == 0 and K <= 104 run on the
    tensor cores (tiles 32, 64, 104), the others on the CUDA-core chunk kernel."""
    ll, mask, pi, u = _hmm_case(100 + K, N=5, T=2400, K=K, ragged=True)
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    z, lg, info = G.sample_hmm_stateseqs(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), chunk_len=300,
                                         burn_in=256, strict=False, return_info=True)
    z, lg = N_(z), N_(lg)
    assert float(info['boundary_err']) < 1e-5, float(info['boundary_err'])
    assert np.all(np.abs(lg - lo) <= 1e-5 * np.abs(lo)), (lg, lo)
    valid = mask > 0
    assert (z[valid] == zo[valid]).mean() >= 0.999
    assert np.all(z[~valid] == 0)


def test_hmm_chunked_rejects_interior_mask(G):
    """The chunked filter assumes trailing padding; a hole in the mask must be flagged (and strict
    mode must then fall back to the sequential kernels, which handle any mask)."""
    ll, mask, pi, u = _hmm_case(5, N=2, T=1500, K=20)
    mask = mask.copy()
    mask[1, 400:420] = 0
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    kw = dict(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), return_info=True)
    _, _, info = G.sample_hmm_stateseqs(chunk_len=300, burn_in=128, strict=False, **kw)
    assert float(info['boundary_err']) > 1.0
    z, lg, info = G.sample_hmm_stateseqs(chunk_len=300, burn_in=128, strict=True, **kw)
    assert float(info['boundary_err']) == 0.0
    valid = mask > 0
    assert (N_(z)[valid] == zo[valid]).mean() >= 0.999
    assert np.all(np.abs(N_(lg) - lo) <= 1e-5 * np.abs(lo))


@pytest.mark.parametrize("D,L,K,T", [(2, 2, 5, 300), (5, 3, 7, 700), (10, 2, 12, 900), (10, 3, 100, 1200)])
def test_ar_stats_empty_states_and_ragged(G, D, L, K, T):
    """Counting sort + tensor-core SYRK with states that never occur and ragged masks."""
    d, s, p, _ = H.problem(seed=300 + D, N=3, T=T, k=4, dk=2, D=D, L=L, K=K, ragged=True)
    z = s['z'].copy()
    z[z == 1] = 0            # state 1 (and any state the simulation skipped) stays empty
    want = O.ar_sufficient_stats(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=z)
    got = N_(G.ar_sufficient_stats(nlags=L, num_states=K, mask=T_(d['mask']), x=T_(s['x']), z=T_(z)))
    scale = np.abs(want).max() + 1e-300
    assert np.abs(got - want).max() / scale < 1e-12
    assert np.all(got[1] == 0.0)
    # run to run bit-reproducible
    again = N_(G.ar_sufficient_stats(nlags=L, num_states=K, mask=T_(d['mask']), x=T_(s['x']), z=T_(z)))
    assert np.array_equal(got, again)


def test_hmm_backward_chunks_are_repaired(G, options):
    """With no warm-up at all every sampler chunk starts from the wrong carry; the device-side
    repair pass must then reproduce, bit for bit, what well warmed-up chunks give."""
    import torch
    ll, mask, pi, u = _hmm_case(21, N=6, T=3000, K=100, ragged=True)
    kw = dict(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), chunk_len=300, burn_in=256, strict=False,
              return_info=True)
    z_ref, l_ref, i_ref = G.sample_hmm_stateseqs(**kw)
    options("hmm_bwd_burn", 0)
    z_rep, l_rep, i_rep = G.sample_hmm_stateseqs(**kw)
    options("hmm_bwd_burn", 3)
    z_rep3, _, _ = G.sample_hmm_stateseqs(**kw)
    torch.cuda.synchronize()
    assert torch.equal(z_ref, z_rep) and torch.equal(z_ref, z_rep3)
    assert torch.equal(l_ref, l_rep)
    zo, lo = O._hmm_batch(ll, mask, pi, u)
    valid = mask > 0
    assert (N_(z_rep)[valid] == zo[valid]).mean() >= 0.999


def test_hmm_forward_chunks_are_repaired_large_K(G, options):
    """K > 128: a chunk whose warmed-up start does not match the carry of its predecessor is filtered
    again on the device from that carry (one round per chunk of a cascade; four chunks per session here, so three rounds).  With a hopeless
    warm-up of 2 frames every boundary fails; after the repair the chunked pass equals the sequential one."""
    import torch
    ll, mask, pi, u = _hmm_case(41, N=2, T=1200, K=200, ragged=True)
    kw = dict(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), return_info=True)
    z0, l0, _ = G.sample_hmm_stateseqs(chunk_len=0, **kw)
    z1, l1, i1 = G.sample_hmm_stateseqs(chunk_len=300, burn_in=2, strict=False, **kw)
    torch.cuda.synchronize()
    assert float(i1['boundary_err']) < 1e-6
    assert float((z0 != z1).float().mean()) < 1e-3
    assert np.all(np.abs(N_(l0) - N_(l1)) <= 1e-5 * np.abs(N_(l0)))
    options("hmm_repair_rounds", 0)
    _, _, i2 = G.sample_hmm_stateseqs(chunk_len=300, burn_in=2, strict=False, **kw)
    assert float(i2['boundary_err']) > 1e-3          # without the repair the mismatch is there and is reported


def test_alternative_kernel_paths_agree(G, options):
    """The kernels kept behind library options stay correct: kalman_impl=0 (CTA-per-chunk
    shared-memory Kalman filter), kalman_impl=1 (lane-per-column warp filter), kalman_obs_mma=0
    (thread-per-frame observation records), emission_pipe=0 and hmm_tc=0 (CUDA-core chunked HMM filter)."""
    import torch
    d, s, p, Gm = H.problem(seed=61, N=3, T=700, k=8, dk=2, D=10, L=3, K=8, ragged=True)
    rng = np.random.default_rng(3)
    w, wl = H.f32r(rng.standard_normal((3, 697, 30))), H.f32r(rng.standard_normal((3, 30)))
    kw = dict(Y=T_(d['Y']), mask=T_(d['mask']), v=T_(s['v']), h=T_(s['h']), z=T_(s['z']), s=T_(s['s']),
              Cd=T_(p['Cd']), sigmasq=T_(p['sigmasq']), Ab=T_(p['Ab']), Q=T_(p['Q']), Gamma=T_(Gm),
              draws={'w': T_(w), 'w_last': T_(wl)}, chunk_len=128, burn_in=160, strict=False)
    x0, i0 = G.resample_latents(0, return_info=True, **kw)
    for name, value in (("kalman_impl", 0), ("kalman_impl", 1), ("kalman_obs_mma", 0)):
        options(name, value)
        x1, i1 = G.resample_latents(0, return_info=True, **kw)
        options(name, 2 if name == "kalman_impl" else 1)
        assert float((x0 - x1).abs().max()) < 1e-5 * max(1.0, float(x0.abs().max())), name
        l0, l1 = N_(i0['loglik']), N_(i1['loglik'])
        assert np.all(np.abs(l0 - l1) <= 1e-9 * np.abs(l0)), (name, l0, l1)
    # emission: the software-pipelined kernel (several blocks per CTA) against the one-block-at-a-time kernel
    d2, s2, p2, _ = H.problem(seed=62, N=2, T=9000, k=4, dk=2, D=10, L=3, K=100, ragged=False)
    e1 = G.ar_log_likelihood(x=T_(s2['x']), Ab=T_(p2['Ab']), Q=T_(p2['Q']))
    options("emission_pipe", 0)
    e0 = G.ar_log_likelihood(x=T_(s2['x']), Ab=T_(p2['Ab']), Q=T_(p2['Q']))
    options("emission_pipe", 1)
    assert torch.equal(e0, e1)
    want = O.ar_log_likelihood(s2['x'], p2['Ab'], p2['Q'])
    err = np.abs(N_(e1).astype(np.float64) - want) / np.maximum(1.0, np.abs(want))
    assert err.max() < 5e-5, err.max()   # 1.8 M float32 entries: the tail of the rounding distribution
    ll, mask, pi, u = _hmm_case(31, N=4, T=2400, K=40, ragged=True)
    hk = dict(ll=T_(ll), mask=T_(mask), pi=T_(pi), u=T_(u), chunk_len=300, burn_in=256, strict=False, return_info=True)
    z0, l0, _ = G.sample_hmm_stateseqs(**hk)
    options("hmm_tc", 0)
    z1, l1, i1 = G.sample_hmm_stateseqs(**hk)
    torch.cuda.synchronize()
    assert float(i1['boundary_err']) < 1e-5
    assert float((z0 != z1).float().mean()) < 1e-3
    assert np.all(np.abs(N_(l0) - N_(l1)) <= 1e-5 * np.abs(N_(l0)))
