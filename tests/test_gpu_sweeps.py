"""GPU tests at the shapes and in the modes the benchmark actually runs: default (auto) chunking
at T = 36 000 against the oracle, a slowly forgetting HMM that must drive the warm-up ladder,
K = 500, sharding invariance of the draws, and the sweep / fit drivers (the cell sequence of
examples/keypoint_slds.ipynb and examples/depth_ar_hmm.ipynb through the `keypoint_moseq` names).

Tolerances (BASELINE.json north_star): marginal log-likelihoods 1e-5 relative, state sequences
>= 99.9 % agreement, sampled latents 2e-6 (float32 storage)."""
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


@pytest.fixture(scope="module")
def long_problem():
    """N=2 sessions x 36 000 frames, k=16, D=10, L=3, K=100 (cfg4's per-session shape), one ragged."""
    N, T, L = 2, 36000, 3
    d, s, p, Gm = H.problem(seed=41, N=N, T=T, k=16, dk=2, D=10, L=L, K=100, ragged=True)
    rng = np.random.default_rng(1)
    dr = dict(u=H.f32r(rng.random((N, T - L))), w=H.f32r(rng.standard_normal((N, T - L, 30))),
              wl=H.f32r(rng.standard_normal((N, 30))))
    return d, s, p, Gm, dr


def _reset_adapt(G):
    G._HMM_ADAPT.clear()
    G._HMM_ADAPT.update(boost=1, streak=0, need=8, probing=False)
    G._LAT_ADAPT.clear()
    G._LAT_ADAPT.update(boost=1, streak=0, need=3, probing=False)
    for a in list(G.FALLBACKS):
        G.FALLBACKS[a] = [] if isinstance(G.FALLBACKS[a], list) else 0


# ------------------------------------------------------------------ default chunking vs the oracle
def test_stateseqs_auto_chunked_T36k_vs_oracle(G, long_problem):
    d, s, p, Gm, dr = long_problem
    zo, lo = O.resample_stateseqs(x=s['x'], mask=d['mask'], Ab=p['Ab'], Q=p['Q'], pi=p['pi'], u=dr['u'])
    _reset_adapt(G)
    import torch
    cl = G.auto_hmm_chunk_len(2, 36000 - 3, 100, 256, torch.device("cuda"))
    assert cl > 0, "this test is about the chunked default"
    z, ll, info = G.resample_stateseqs(0, x=T_(s['x']), mask=T_(d['mask']), Ab=T_(p['Ab']), Q=T_(p['Q']),
                                       pi=T_(p['pi']), draws={'u': T_(dr['u'])}, return_info=True)
    assert info["chunk_len"] == cl
    assert G.FALLBACKS["stateseqs"] == 0 and float(info["boundary_err"]) <= G.STATESEQ_CHUNKING["tol"]
    valid = d['mask'][:, 3:] > 0
    assert (N_(z)[valid] == zo[valid]).mean() >= 0.999
    assert np.all(N_(z)[~valid] == 0)
    assert np.all(np.abs(N_(ll) - lo) <= 1e-5 * np.abs(lo)), (N_(ll), lo)


@pytest.mark.parametrize("chunk", ["auto", 2572])
def test_latents_chunked_T36k_vs_oracle(G, long_problem, chunk):
    """The default ("auto") chunking and the chunk length the benchmark shard ends up with
    (125 sessions: 2572 steps, 256 warm-up) against the sequential float64 oracle."""
    d, s, p, Gm, dr = long_problem
    xo, llo = O.resample_latents(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], z=s['z'], s=s['s'], Cd=p['Cd'],
                                 sigmasq=p['sigmasq'], Ab=p['Ab'], Q=p['Q'], Gamma=Gm, w=dr['w'], w_last=dr['wl'],
                                 return_loglik=True)
    _reset_adapt(G)
    x, info = G.resample_latents(0, Y=T_(d['Y']), mask=T_(d['mask']), v=T_(s['v']), h=T_(s['h']), z=T_(s['z']),
                                 s=T_(s['s']), Cd=T_(p['Cd']), sigmasq=T_(p['sigmasq']), Ab=T_(p['Ab']), Q=T_(p['Q']),
                                 Gamma=T_(Gm), draws={'w': T_(dr['w']), 'w_last': T_(dr['wl'])}, return_info=True,
                                 chunk_len=chunk, burn_in=256)
    assert info["chunk_len"] > 0 and G.FALLBACKS["latents"] == 0
    assert N_(info["boundary_err"]).max() <= G.LATENTS_CHUNKING["tol"]
    x = N_(x).astype(np.float64)
    assert np.abs(x - xo).max() < 2e-6 * max(1.0, np.abs(xo).max()), np.abs(x - xo).max()
    assert np.all(np.abs(N_(info["loglik"]) - llo) <= 1e-9 * np.abs(llo))


# ------------------------------------------------------------------ slowly forgetting HMM
def _slow_case(seed, N, T, K, stick):
    """States that are hardly distinguishable (all Ab_k within 1 % of each other, the regime right
    after initial_ar_params) and very sticky transitions: the filter forgets its start slowly."""
    d, s, p, Gm = H.problem(seed=seed, N=N, T=T, k=8, dk=2, D=10, L=3, K=K, ragged=True)
    rng = np.random.default_rng(seed)
    Ab = np.repeat(p['Ab'][:1], K, axis=0) * (1 + 0.01 * rng.standard_normal((K, 1, 1)))
    Q = np.repeat(p['Q'][:1], K, axis=0)
    pi = (1 - stick) * rng.dirichlet(np.ones(K), size=K) + stick * np.eye(K)
    u = H.f32r(rng.random((N, T - 3)))
    return d, s, Ab, Q, pi, u


@pytest.mark.parametrize("stick,max_boost,expect_exact", [(0.98, 16, False), (0.9995, 16, False), (0.9995, 2, True)])
def test_slow_forgetting_hmm_drives_the_retry_ladder(G, monkeypatch, stick, max_boost, expect_exact):
    """(0.9995, 16): warm-ups of 32, 64, 128, 256 frames all fail, 512 passes; with the ladder capped
    at 2x the sequential kernels have to take over."""
    monkeypatch.setattr(G, "HMM_MAX_BOOST", max_boost)
    N, T, K = 3, 9000, 20
    d, s, Ab, Q, pi, u = _slow_case(7, N, T, K, stick)
    zo, lo = O.resample_stateseqs(x=s['x'], mask=d['mask'], Ab=Ab, Q=Q, pi=pi, u=u)
    kw = dict(x=T_(s['x']), mask=T_(d['mask']), Ab=T_(Ab), Q=T_(Q), pi=T_(pi), draws={'u': T_(u)})
    # reference point on the device: the sequential kernels
    z_seq, ll_seq = G.resample_stateseqs(0, **kw, chunk_len=0)
    # short chunks with a warm-up that is too short for this model: strict mode must notice and
    # lengthen it (or give up and run sequentially) until the tolerances hold
    _reset_adapt(G)
    z, ll, info = G.resample_stateseqs(0, **kw, chunk_len=300, burn_in=32, strict=True, return_info=True)
    assert G.FALLBACKS["stateseqs"] >= 1, "the warm-up ladder was not exercised"
    assert (G.FALLBACKS.get("stateseqs_exact", 0) >= 1) == expect_exact
    # without verification the same configuration is visibly off
    _, _, raw = G.resample_stateseqs(0, **kw, chunk_len=300, burn_in=32, strict=False, return_info=True)
    assert float(raw["boundary_err"]) > G.STATESEQ_CHUNKING["tol"]
    valid = d['mask'][:, 3:] > 0
    assert np.all(np.abs(N_(ll) - lo) <= 1e-5 * np.abs(lo)), (N_(ll), lo)
    assert np.all(np.abs(N_(ll_seq) - lo) <= 1e-5 * np.abs(lo))
    assert (N_(z)[valid] == N_(z_seq)[valid]).mean() >= 0.999
    # against the float64 oracle, flat posteriors make float32 ties more frequent; still within contract
    assert (N_(z_seq)[valid] == zo[valid]).mean() >= 0.999
    assert (N_(z)[valid] == zo[valid]).mean() >= 0.999
    _reset_adapt(G)


# ------------------------------------------------------------------ K = 500 (cfg5)
def test_stateseqs_K500_parity(G):
    N, T, K, L = 3, 1200, 500, 3
    d, s, p, Gm = H.problem(seed=55, N=N, T=T, k=8, dk=2, D=10, L=L, K=K, ragged=True)
    u = H.f32r(np.random.default_rng(5).random((N, T - L)))
    zo, lo = O.resample_stateseqs(x=s['x'], mask=d['mask'], Ab=p['Ab'], Q=p['Q'], pi=p['pi'], u=u)
    z, ll = G.resample_stateseqs(0, x=T_(s['x']), mask=T_(d['mask']), Ab=T_(p['Ab']), Q=T_(p['Q']), pi=T_(p['pi']),
                                 draws={'u': T_(u)})
    valid = d['mask'][:, L:] > 0
    assert (N_(z)[valid] == zo[valid]).mean() >= 0.999
    assert np.all(np.abs(N_(ll) - lo) <= 1e-5 * np.abs(lo))
    want = O.ar_log_likelihood(s['x'], p['Ab'], p['Q'])
    got = N_(G.ar_log_likelihood(x=T_(s['x']), Ab=T_(p['Ab']), Q=T_(p['Q']))).astype(np.float64)
    assert (np.abs(got - want) / np.maximum(1.0, np.abs(want))).max() < 5e-5     # 3xTF32, 1.8 M entries
    F = N_(G.ar_sufficient_stats(nlags=L, num_states=K, mask=T_(d['mask']), x=T_(s['x']), z=T_(s['z'])))
    Fo = O.ar_sufficient_stats(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=s['z'])
    assert np.abs(F - Fo).max() <= 1e-10 * np.abs(Fo).max()


# ------------------------------------------------------------------ sharding invariance of the draws
def test_session_draws_do_not_depend_on_sharding(G):
    """Session n's Philox draws are keyed by its GLOBAL index: running sessions [lo, hi) as a shard
    with session_offset = lo reproduces, bit for bit, rows lo..hi of the unsharded call."""
    import torch
    N, T, L, K = 6, 500, 3, 8
    d, s, p, Gm = H.problem(seed=63, N=N, T=T, k=8, dk=2, D=10, L=L, K=K, ragged=True)
    hp = H.hyp(10, L, K)
    dd, ss, pp = H.to_torch(d), H.to_torch(s), H.to_torch(p)
    dd = {a: b.float() for a, b in dd.items()}
    ss = {a: (b if a == 'z' else b.float()) for a, b in ss.items()}
    key = (5, 17)

    def steps(lo, hi):
        sl = slice(lo, hi)
        D_ = {a: b[sl].contiguous() for a, b in dd.items()}
        S_ = {a: b[sl].contiguous() for a, b in ss.items()}
        kw = dict(Gamma=T_(Gm), session_offset=lo)
        out = {}
        out['z'] = G.resample_stateseqs(key, **D_, **S_, **pp, chunk_len=0, session_offset=lo)[0]
        out['x'] = G.resample_latents(key, **D_, **S_, **pp, chunk_len=0, **kw)
        out['h'] = G.resample_heading(key, **D_, **S_, **pp, **kw)
        out['v'] = G.resample_location(key, **D_, **S_, **pp, sigmasq_loc=hp['sigmasq_loc'], **kw)
        out['s'] = G.resample_scales(key, **D_, **S_, **pp, nu_s=hp['nu_s'], s_0=hp['s_0'], **kw)
        return out

    full = steps(0, N)
    for lo, hi in [(0, 2), (2, 6), (5, 6)]:
        part = steps(lo, hi)
        for name in full:
            assert torch.equal(part[name], full[name][lo:hi]), (name, lo, hi)
    # and a different offset gives different draws
    other = G.resample_scales(key, **{a: b[0:2].contiguous() for a, b in dd.items()},
                              **{a: b[0:2].contiguous() for a, b in ss.items()}, **pp, nu_s=hp['nu_s'], s_0=hp['s_0'],
                              Gamma=T_(Gm), session_offset=3)
    assert not torch.equal(other, full['s'][0:2])


def test_fill_at_offsets_tile_the_global_stream():
    import torch
    from keypointmoseq_b200 import distributions as dist
    for fill in (dist.fill_uniform, dist.fill_normal):
        whole = fill((10, 7, 3), 99, 4, "cuda")
        for lo, hi in [(0, 3), (3, 4), (4, 10)]:
            assert torch.equal(fill((hi - lo, 7, 3), 99, 4, "cuda", first_row=lo), whole[lo:hi])
    whole = dist.fill_gamma((9, 5), 4.0, 99, 4, "cuda")
    assert torch.equal(dist.fill_gamma((4, 5), 4.0, 99, 4, "cuda", first_row=5), whole[5:9])


# ------------------------------------------------------------------ regression utility
def test_resample_regression_params_equals_ar_params(G):
    """gibbs.py:127-136 with explicit regressors for one state = resample_ar_params restricted to it."""
    import torch
    from keypointmoseq_b200 import util as U
    D, L, K = 10, 3, 4
    d, s, p, Gm = H.problem(seed=23, N=2, T=300, D=D, L=L, K=K, ragged=True)
    hp = H.hyp(D, L, K)
    rng = np.random.default_rng(2)
    dr = {"chi2": rng.chisquare(hp['nu_0'] + 40.0, size=(K, D)), "normals": rng.standard_normal((K, D * (D - 1) // 2)),
          "G": rng.standard_normal((K, D, D * L + 1))}
    pri = dict(nu_0=hp['nu_0'], S_0=T_(hp['S_0']), M_0=T_(hp['M_0']), K_0=T_(hp['K_0']))
    Ab, Q = G.resample_ar_params(0, nlags=L, num_states=K, mask=T_(d['mask']), x=T_(s['x']), z=T_(s['z']), **pri,
                                 draws={a: T_(b) for a, b in dr.items()})
    x = T_(s['x'])
    x_in = U.pad_affine(U.get_lags(x, L)).reshape(-1, D * L + 1)
    x_out = x[:, L:].reshape(-1, D)
    for kk in (0, 2):
        m = (T_(d['mask'])[:, L:].reshape(-1) * (T_(s['z']).reshape(-1) == kk)).double()
        Ab1, Q1 = G.resample_regression_params(0, m, x_in, x_out, **pri,
                                               draws={a: T_(b[kk:kk + 1]) for a, b in dr.items()})
        assert torch.allclose(Ab1, Ab[kk], rtol=1e-8, atol=1e-10) and torch.allclose(Q1, Q[kk], rtol=1e-8, atol=1e-12)
    Abo, Qo = O.resample_ar_params(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=s['z'], nu_0=hp['nu_0'],
                                   S_0=hp['S_0'], M_0=hp['M_0'], K_0=hp['K_0'], chi2=dr['chi2'], normals=dr['normals'],
                                   G=dr['G'])
    assert np.allclose(N_(Ab), Abo, rtol=1e-6, atol=1e-9) and np.allclose(N_(Q), Qo, rtol=1e-7, atol=1e-12)


# ------------------------------------------------------------------ sweeps, fit, notebook sequence
def _device_problem(seed, N, T, k, D, L, K):
    d, s, p, Gm = H.problem(seed=seed, N=N, T=T, k=k, dk=2, D=D, L=L, K=K, ragged=True)
    dd = {a: T_(b).float() for a, b in d.items()}
    ss = {a: (T_(b.astype(np.int32)) if a == 'z' else T_(b).float()) for a, b in s.items()}
    pp = {a: T_(b) for a, b in p.items()}
    return d, s, p, Gm, dd, ss, pp


def test_gibbs_sweep_equals_step_sequence(G):
    """parallel.gibbs_sweep (optimistic launches, one verification read-back) = the eight steps of
    keypoint_slds.ipynb cell 12 called one after the other with the same key."""
    import torch
    from keypointmoseq_b200 import parallel
    N, T, k, D, L, K = 3, 400, 8, 10, 3, 6
    d, s, p, Gm, dd, ss, pp = _device_problem(71, N, T, k, D, L, K)
    hp = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in H.hyp(D, L, K).items()}
    key = (3, 4)
    st, pr = parallel.gibbs_sweep(key, dd, ss, pp, hp, Gamma=T_(Gm))
    kw = dict(Gamma=T_(Gm))
    q, t = dict(pp), dict(ss)
    q['Ab'], q['Q'] = G.resample_ar_params(key, **dd, **t, **q, **hp)
    q['sigmasq'] = G.resample_obs_variance(key, **dd, **t, **q, **hp, **kw)
    q['betas'], q['pi'] = G.resample_hdp_transitions(key, **dd, **t, **q, **hp)
    t['z'] = G.resample_stateseqs(key, **dd, **t, **q)[0]
    t['x'] = G.resample_latents(key, **dd, **t, **q, **kw)
    t['h'] = G.resample_heading(key, **dd, **t, **q, **kw)
    t['v'] = G.resample_location(key, **dd, **t, **q, **hp, **kw)
    t['s'] = G.resample_scales(key, **dd, **t, **q, **hp, **kw)
    for name in ('Ab', 'Q', 'sigmasq', 'betas', 'pi'):
        assert torch.equal(pr[name], q[name]), name
    for name in ('z', 'x', 'h', 'v', 's'):
        assert torch.equal(st[name], t[name]), name
    assert set(ss) == {a for a in st if not a.startswith("_")} and all(ss[a] is not st[a] for a in ('z', 'x', 'h', 'v', 's'))   # inputs untouched


def test_gibbs_sweep_reuses_scales_pass_statistics(G):
    """The scales step leaves next sweep's observation-noise statistics behind (same pass over Y): a
    second sweep that gets the states back untouched gives exactly what it gives when it recomputes
    them, and the statistics themselves equal kpms_obs_variance_stats on the new state."""
    import torch
    from keypointmoseq_b200 import parallel
    N, T, k, D, L, K = 3, 400, 8, 10, 3, 6
    d, s, p, Gm, dd, ss, pp = _device_problem(73, N, T, k, D, L, K)
    hp = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in H.hyp(D, L, K).items()}
    st1, pr1 = parallel.gibbs_sweep((2, 0), dd, ss, pp, hp, Gamma=T_(Gm))
    cache = st1[parallel.OBS_STATS_KEY]
    S_ref, m_ref = G.obs_variance_stats(Y=dd["Y"], mask=dd["mask"], Cd=pr1["Cd"], v=st1["v"], h=st1["h"], x=st1["x"],
                                        s=st1["s"], Gamma=T_(Gm))
    assert torch.allclose(cache["S_y"], S_ref, rtol=1e-12) and torch.equal(cache["msum"], m_ref)
    st2a, pr2a = parallel.gibbs_sweep((2, 1), dd, st1, pr1, hp, Gamma=T_(Gm))
    plain = {a: b for a, b in st1.items() if not a.startswith("_")}
    st2b, pr2b = parallel.gibbs_sweep((2, 1), dd, plain, pr1, hp, Gamma=T_(Gm))
    assert torch.allclose(pr2a["sigmasq"], pr2b["sigmasq"], rtol=1e-12)
    for name in ("z", "x", "h", "v", "s"):
        assert torch.equal(st2a[name], st2b[name]) or torch.allclose(st2a[name].double(), st2b[name].double(), rtol=1e-5), name
    # a state modified in place invalidates the cache
    st1["s"].mul_(1.5)
    st3, pr3 = parallel.gibbs_sweep((2, 1), dd, st1, pr1, hp, Gamma=T_(Gm))
    S_mod, m_mod = G.obs_variance_stats(Y=dd["Y"], mask=dd["mask"], Cd=pr1["Cd"], v=st1["v"], h=st1["h"], x=st1["x"],
                                        s=st1["s"], Gamma=T_(Gm))
    want = G.sample_obs_variance((2, 1), S_mod, m_mod, nu_sigma=hp["nu_sigma"], sigmasq_0=hp["sigmasq_0"])
    assert torch.allclose(pr3["sigmasq"], want, rtol=1e-12)


def test_gibbs_sweep_with_parameters_drawn_ahead(G):
    """next_key: the following sweep's AR / transition parameters are drawn at the end of this one (off
    the critical path).  Three chained sweeps give bit-for-bit what three plain sweeps give; a caller who
    comes back with another key or other tensors simply gets them recomputed."""
    import torch
    from keypointmoseq_b200 import parallel
    N, T, k, D, L, K = 3, 400, 8, 10, 3, 6
    d, s, p, Gm, dd, ss, pp = _device_problem(75, N, T, k, D, L, K)
    hp = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in H.hyp(D, L, K).items()}
    sa, pa = ss, pp
    sb, pb = ss, pp
    for i in range(3):
        sa, pa = parallel.gibbs_sweep((4, i), dd, sa, pa, hp, Gamma=T_(Gm), next_key=(4, i + 1))
        assert parallel.NEXT_PARAMS_KEY in sa
        plain = {a: b for a, b in sb.items() if not a.startswith("_")}
        sb, pb = parallel.gibbs_sweep((4, i), dd, plain, pb, hp, Gamma=T_(Gm))
    for name in ("z", "x", "h", "v", "s"):
        assert torch.equal(sa[name], sb[name]), name
    for name in ("Ab", "Q", "pi", "betas", "sigmasq"):
        assert torch.equal(pa[name], pb[name]), name
    # another key than the one announced: the cache is ignored
    sc, pc = parallel.gibbs_sweep((9, 9), dd, sa, pa, hp, Gamma=T_(Gm))
    plain = {a: b for a, b in sa.items() if not a.startswith("_")}
    sd, pd_ = parallel.gibbs_sweep((9, 9), dd, plain, pa, hp, Gamma=T_(Gm))
    assert torch.equal(pc["Ab"], pd_["Ab"]) and torch.equal(sc["z"], sd["z"])


def test_heading_and_scales_in_kernel_draws(G):
    """Default (no injected draws) paths: scales with in-kernel Gamma variates equal the injected path fed
    with kpms_fill_gamma's values; the heading sampler never exhausts its proposals and follows the von Mises
    concentration."""
    import torch
    from keypointmoseq_b200 import distributions as dist
    N, T, k, D, L, K = 3, 300, 8, 10, 3, 5
    d, s, p, Gm, dd, ss, pp = _device_problem(74, N, T, k, D, L, K)
    hp = H.hyp(D, L, K)
    kw = dict(**dd, **ss, **pp, Gamma=T_(Gm), nu_s=hp["nu_s"], s_0=hp["s_0"])
    s_a = G.resample_scales(9, **kw, session_offset=2)
    g = dist.fill_gamma((N, T, k), (hp["nu_s"] + 3) / 2, dist.seed_of(9), dist.STREAM_S, "cuda", first_row=2)
    s_b = G.resample_scales(9, **kw, draws={"g": g})
    assert torch.equal(s_a, s_b)
    h, info = G.resample_heading(9, **dd, **ss, **pp, Gamma=T_(Gm), return_info=True)
    assert int(info["exhausted"]) == 0
    # first-moment test over all frames: E cos(h - theta) = I1(kappa) / I0(kappa), Var = 1 - A / kappa - A^2
    dev_ = h.double() - info["theta"]
    kap = info["kappa"].clamp(min=1e-6)
    A = torch.special.i1e(kap) / torch.special.i0e(kap)
    var = (1.0 - A / kap - A * A).clamp(min=1e-12)
    zscore = float((torch.cos(dev_) - A).sum() / var.sum().sqrt())
    assert abs(zscore) < 4.5, zscore
    assert abs(float(torch.sin(dev_).sum() / (A / kap).clamp(min=1e-12).sum().sqrt())) < 4.5   # symmetric about theta
    h2 = G.resample_heading(9, **dd, **ss, **pp, Gamma=T_(Gm))
    assert torch.equal(h, h2)


def test_gibbs_sweep_repairs_a_failed_chunk_check(G):
    """A sweep whose optimistic chunked passes fail verification redoes them: the result equals the
    sweep run with the sequential kernels."""
    import torch
    from keypointmoseq_b200 import parallel
    N, T, k, D, L, K = 2, 2400, 8, 10, 3, 6
    d, s, p, Gm, dd, ss, pp = _device_problem(72, N, T, k, D, L, K)
    hp = {a: (T_(b) if isinstance(b, np.ndarray) else b) for a, b in H.hyp(D, L, K).items()}
    saved_l, saved_s = dict(G.LATENTS_CHUNKING), dict(G.STATESEQ_CHUNKING)
    try:
        G.LATENTS_CHUNKING.update(chunk_len=0)
        G.STATESEQ_CHUNKING.update(chunk_len=0)
        st0, pr0 = parallel.gibbs_sweep((1, 1), dd, ss, pp, hp, Gamma=T_(Gm))
        _reset_adapt(G)
        G.LATENTS_CHUNKING.update(chunk_len=300, burn_in=2)       # hopeless warm-up: must be caught and redone
        st1, pr1 = parallel.gibbs_sweep((1, 1), dd, ss, pp, hp, Gamma=T_(Gm))
        assert G.FALLBACKS["latents"] >= 1
        assert G._LAT_ADAPT["boost"] == 1          # a lone failure is redone and costs nothing more
        for name in ('z', 'x', 'h', 'v', 's'):
            assert torch.equal(st0[name], st1[name]), name
        parallel.gibbs_sweep((1, 2), dd, st1, pr1, hp, Gamma=T_(Gm))
        assert G._LAT_ADAPT["boost"] >= 2          # failing again right away: the next sweeps start from a longer warm-up
    finally:
        G.LATENTS_CHUNKING.clear()
        G.LATENTS_CHUNKING.update(saved_l)
        G.STATESEQ_CHUNKING.clear()
        G.STATESEQ_CHUNKING.update(saved_s)
        _reset_adapt(G)


def test_notebook_cell_sequence_through_keypoint_moseq_names():
    """examples/keypoint_slds.ipynb cells 2-12 with `keypoint_moseq` imports and torch CUDA tensors in
    place of JAX arrays: ingest, initialisation, AR-only warm-up, full sweeps."""
    import torch
    from keypoint_moseq.util import expected_keypoints, merge_data, stateseq_stats
    from keypoint_moseq.gibbs import (resample_ar_params, resample_hdp_transitions, resample_heading,
                                      resample_latents, resample_location, resample_obs_variance,
                                      resample_scales, resample_stateseqs)
    from keypoint_moseq.initialize import (initial_ar_params, initial_hdp_transitions, initial_heading,
                                           initial_latents, initial_location)
    from keypointmoseq_b200.synthetic import simulate
    latent_dim, num_states, nlags = 10, 20, 3
    raw, _, _, _ = simulate(81, 3, 700, k=8, dk=2, D=latent_dim, L=nlags, K=6, ragged=False, dtype=np.float32)
    lens = (700, 520, 610)
    keypoint_data_dict = {"sess%d" % i: raw["Y"][i, :n] for i, n in enumerate(lens)}            # cell 2
    Y, mask, keys = merge_data(keypoint_data_dict, device="cuda")
    assert Y.shape == (3, 700, 8, 2) and mask.sum() == sum(lens) and len(keys) == 3
    eye = lambda n: torch.eye(n, dtype=torch.float64, device="cuda")
    trans_hypparams = {'gamma': 1e3, 'alpha': 5.7, 'kappa': 1e6, 'num_states': num_states}     # cell 4
    ar_hypparams = {'nu_0': latent_dim + 2, 'S_0': .01 * eye(latent_dim),
                    'M_0': torch.nn.functional.pad(eye(latent_dim), ((nlags - 1) * latent_dim, 1)),
                    'K_0': 10 * eye(latent_dim * nlags + 1), 'num_states': num_states, 'nlags': nlags}
    obs_hypparams = {'sigmasq_0': 10, 'sigmasq_C': .1, 'nu_sigma': 1e5, 'nu_s': 5, 's_0': 1}
    translation_hypparams = {'sigmasq_loc': 0.5}
    key = 0                                                                                    # cell 6
    data = {'mask': mask, 'Y': Y}
    states, params = {}, {}
    states['v'] = initial_location(**data)
    states['h'] = initial_heading(torch.tensor([0, 1, 2]), torch.tensor([5, 6, 7]), **data)
    states['x'], params['Cd'], pca_model = initial_latents(latent_dim=latent_dim, **data, **states, num_samples=1500)
    assert len(pca_model.explained_variance_ratio_) == latent_dim
    params['betas'], params['pi'] = initial_hdp_transitions(key, **trans_hypparams)
    params['Ab'], params['Q'] = initial_ar_params(key, **ar_hypparams, device="cuda")
    params['sigmasq'] = torch.ones(Y.shape[-2], dtype=torch.float64, device="cuda")
    states['z'], _ = resample_stateseqs(key, **data, **states, **params)
    states['s'] = resample_scales(key, **data, **states, **params, **obs_hypparams)
    for i in range(3):                                                                         # cell 9
        params['betas'], params['pi'] = resample_hdp_transitions((key, i), **data, **states, **params, **trans_hypparams)
        params['Ab'], params['Q'] = resample_ar_params((key, i), **data, **states, **params, **ar_hypparams)
        states['z'], _ = resample_stateseqs((key, i), **data, **states, **params)
    usage, durations = stateseq_stats(states['z'], mask)
    assert int(usage.sum()) == int(mask[:, nlags:].sum()) and int(durations.sum()) == int(usage.sum())
    trans_hypparams = {'gamma': 1e3, 'alpha': 100, 'kappa': 1e6 / 50, 'num_states': num_states}  # cell 11
    for i in range(3):                                                                         # cell 12
        params['Ab'], params['Q'] = resample_ar_params((key, i), **data, **states, **params, **ar_hypparams)
        params['sigmasq'] = resample_obs_variance((key, i), **data, **states, **params, **obs_hypparams)
        params['betas'], params['pi'] = resample_hdp_transitions((key, i), **data, **states, **params, **trans_hypparams)
        states['z'] = resample_stateseqs((key, i), **data, **states, **params)[0]
        states['x'] = resample_latents((key, i), **data, **states, **params)
        states['h'] = resample_heading((key, i), **data, **states, **params)
        states['v'] = resample_location(key, **data, **states, **params, **translation_hypparams)
        states['s'] = resample_scales((key, i), **data, **states, **params, **obs_hypparams)
    for name, t in list(states.items()) + list(params.items()):
        assert torch.isfinite(t.double()).all(), name
    assert states['x'].shape == (3, 700, latent_dim) and states['z'].shape == (3, 700 - nlags)
    live = mask > 0
    Yexp = expected_keypoints(Y=Y, v=states['v'], h=states['h'], x=states['x'], Cd=params['Cd']).float()
    assert float((Y - Yexp)[live].pow(2).mean().sqrt()) < 3.0 * float(Y[live].std())   # the fit explains the keypoints


def test_fit_driver_checkpoints_and_resumes(G, tmp_path):
    """fit(): AR-only warm-up then full sweeps; a run interrupted after its checkpoint and resumed gives
    the same states and parameters as the uninterrupted run; the AR-HMM mode (depth_ar_hmm.ipynb) runs."""
    import torch
    from keypointmoseq_b200 import fitting
    N, T, k, D, L, K = 3, 500, 8, 10, 3, 8
    d, s, p, Gm, dd, ss, pp = _device_problem(91, N, T, k, D, L, K)
    from keypointmoseq_b200.synthetic import default_hypparams
    trans, ar, obs, loc = default_hypparams(D, L, K)
    kw = dict(trans_hypparams=trans, ar_hypparams=ar, obs_hypparams=obs, translation_hypparams=loc, Gamma=T_(Gm),
              ar_only_iters=2, ar_only_trans_hypparams=dict(trans, kappa=1e6, alpha=5.7))
    _reset_adapt(G)
    st_a, pr_a, hist = fitting.fit(11, dd, ss, pp, num_iters=3, log_every=1, log_fn=None, **kw)
    assert [h["phase"] for h in hist] == ["ar_only"] * 2 + ["full"] * 3
    ck = str(tmp_path / "ck")
    _reset_adapt(G)
    fitting.fit(11, dd, ss, pp, num_iters=1, checkpoint_dir=ck, checkpoint_every=1, **kw)       # sweeps 0..2
    loaded = fitting.load_checkpoint(ck)
    assert loaded["iteration"] == 2 and set(loaded["states"]) == set(ss)
    _reset_adapt(G)
    st_b, pr_b, _ = fitting.fit(11, dd, ss, pp, num_iters=3, checkpoint_dir=ck, checkpoint_every=10, **kw)   # resumes at 3
    for name in st_a:
        assert torch.equal(st_a[name], st_b[name]), name
    for name in pr_a:
        assert torch.equal(pr_a[name], pr_b[name]), name
    # AR-HMM on given x (depth_ar_hmm.ipynb cells 6-8)
    st_c, pr_c, _ = fitting.fit(5, {"mask": dd["mask"]}, {"x": ss["x"], "z": ss["z"]},
                                {a: pp[a] for a in ("Ab", "Q", "pi", "betas")}, trans_hypparams=trans,
                                ar_hypparams=ar, num_iters=3)
    assert st_c["z"].shape == ss["z"].shape and torch.isfinite(pr_c["Ab"]).all()
    assert torch.equal(st_c["x"], ss["x"])
    _reset_adapt(G)
