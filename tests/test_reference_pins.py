"""Oracle and CUDA kernels against outputs of the REFERENCE'S OWN numba code
(tests/golden/reference_pins.npz, made by tests/golden/make_reference_pins.py in the build
container from /root/reference): rows a10 (count_transitions, util.py:178-189) and a11
(sample_crp_tablecounts, transitions.py:7-21).  Integer work: bit-exact."""
import os

import numpy as np
import pytest

from oracle import kpms_oracle as O

PINS = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_pins.npz"))


def _crp_case(ci):
    customers = PINS[f"crp{ci}_customers"]
    uni = np.random.RandomState(int(PINS[f"crp{ci}_seed"])).random_sample(int(customers.sum()))
    return float(PINS[f"crp{ci}_alpha"]), customers, PINS[f"crp{ci}_betas"], uni, PINS[f"crp{ci}_m"]


@pytest.mark.parametrize("ci", range(int(PINS["n_crp"])))
def test_oracle_crp_tablecounts_equals_reference(ci):
    alpha, customers, betas, uni, m_ref = _crp_case(ci)
    assert np.array_equal(O.sample_crp_tablecounts(alpha, customers, betas, uni), m_ref)


@pytest.mark.parametrize("ci", range(int(PINS["n_cnt"])))
def test_oracle_count_transitions_equals_reference(ci):
    got = O.count_transitions(int(PINS[f"cnt{ci}_K"]), PINS[f"cnt{ci}_mask"], PINS[f"cnt{ci}_z"])
    assert np.array_equal(got, PINS[f"cnt{ci}_counts"])


@pytest.mark.gpu
@pytest.mark.parametrize("ci", range(int(PINS["n_crp"])))
def test_cuda_crp_tablecounts_equals_reference(ci):
    import torch
    from keypointmoseq_b200 import gibbs as G
    alpha, customers, betas, uni, m_ref = _crp_case(ci)
    K = customers.shape[0]
    dev = "cuda"
    draws = {"randseq": torch.as_tensor(uni, device=dev), "u_thin": torch.zeros((K, 4), device=dev)}
    _, ms_raw = G.sample_ms(0, torch.as_tensor(customers, device=dev), torch.as_tensor(betas, device=dev),
                            alpha, 1e6, draws=draws)
    assert np.array_equal(ms_raw.cpu().numpy(), m_ref)


@pytest.mark.gpu
@pytest.mark.parametrize("ci", range(int(PINS["n_cnt"])))
def test_cuda_count_transitions_equals_reference(ci):
    import torch
    from keypointmoseq_b200 import gibbs as G
    K = int(PINS[f"cnt{ci}_K"])
    z = torch.as_tensor(PINS[f"cnt{ci}_z"].astype(np.int32), device="cuda")
    mask = torch.as_tensor(PINS[f"cnt{ci}_mask"].astype(np.float32), device="cuda")
    got = G.transition_counts(z=z, mask=mask, num_states=K)
    assert np.array_equal(got.cpu().numpy().astype(np.float64), PINS[f"cnt{ci}_counts"])


# ====================================================================== the resample_* steps
# tests/golden/reference_sweep.npz: the reference's own gibbs.py / kalman.py / distributions.py /
# util.py executed under a NumPy stand-in for the JAX calls they make (tests/golden/jax_numpy_shim.py),
# with injected draws.  Per-frame inputs and draws are float32-representable, parameters float64.
# reference_sweep_bench.npz: the same at the shapes the benchmark dispatches to (D=10, L=3, k=16), plus
# resample_hdp_transitions.
_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIX = {"small": np.load(os.path.join(_GOLD, "reference_sweep.npz")),
       "bench": np.load(os.path.join(_GOLD, "reference_sweep_bench.npz"))}
both = pytest.mark.parametrize("fx", ["small", "bench"])


def _dims(SW):
    return [int(v) for v in SW["dims"]]


def _kw(SW):
    k_ = _dims(SW)[2]
    return dict(Y=SW["Y"], mask=SW["mask"], x=SW["x"], v=SW["v"], h=SW["h"], s=SW["s"], z=SW["z"], Ab=SW["Ab"],
                Q=SW["Q"], pi=SW["pi"], Cd=SW["Cd"], sigmasq=SW["sigmasq"], Gamma=O.center_embedding(k_))


def _rel(a, b):
    return float(np.abs(np.asarray(a, dtype=np.float64) - b).max() / (np.abs(b).max() + 1e-300))


@both
def test_oracle_stateseqs_equals_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    z, ll = O.resample_stateseqs(x=SW["x"], mask=SW["mask"], Ab=SW["Ab"], Q=SW["Q"], pi=SW["pi"], u=SW["d_stateseq_u"])
    assert np.array_equal(z, SW["r_stateseq_z"]) and _rel(ll, SW["r_stateseq_ll"]) < 1e-12


@both
def test_oracle_latents_equals_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    x = O.resample_latents(**_kw(SW), w=SW["d_latents_w"], w_last=SW["d_latents_w_last"])
    assert _rel(x, SW["r_latents_x"]) < 1e-11


@both
def test_oracle_heading_location_scales_obsvar_equal_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    kw = _kw(SW)
    h = O.resample_heading(**kw, u=SW["d_heading_u"])
    assert np.abs(np.angle(np.exp(1j * (h - SW["r_heading_h"])))).max() < 1e-12
    v = O.resample_location(**kw, sigmasq_loc=float(SW["hyp_sigmasq_loc"]), w=SW["d_location_w"],
                            w_last=SW["d_location_w_last"])
    assert _rel(v, SW["r_location_v"]) < 1e-12
    s = O.resample_scales(**kw, nu_s=float(SW["hyp_nu_s"]), s_0=float(SW["hyp_s_0"]), g=SW["d_scales_g"])
    assert _rel(s, SW["r_scales_s"]) < 1e-12
    sg = O.resample_obs_variance(**kw, nu_sigma=float(SW["hyp_nu_sigma"]), sigmasq_0=float(SW["hyp_sigmasq_0"]),
                                 g=SW["d_obsvar_g"])
    assert _rel(sg, SW["r_obsvar_sigmasq"]) < 1e-12


@both
def test_oracle_ar_params_equal_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    Ab, Q = O.resample_ar_params(nlags=L_, num_states=K_, mask=SW["mask"], x=SW["x"], z=SW["z"],
                                 nu_0=float(SW["hyp_nu_0"]), S_0=SW["hyp_S_0"], M_0=SW["hyp_M_0"], K_0=SW["hyp_K_0"],
                                 chi2=SW["d_ar_chi2"], normals=SW["d_ar_normals"], G=SW["d_ar_G"])
    assert _rel(Ab, SW["r_ar_Ab"]) < 1e-10 and _rel(Q, SW["r_ar_Q"]) < 1e-10


@both
def test_oracle_obs_params_equals_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    Cd = O.resample_obs_params(Y=SW["Y"], mask=SW["mask"], sigmasq=SW["sigmasq"], v=SW["v"], h=SW["h"], x=SW["x"],
                               s=SW["s"], sigmasq_C=float(SW["hyp_sigmasq_C"]), Gamma=O.center_embedding(k_),
                               w=SW["d_obsparams_w"])
    assert _rel(Cd, SW["r_obsparams_Cd"]) < 1e-11


@both
def test_oracle_diagnostics_equal_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    lj = O.log_joint_likelihood(**_kw(SW), sigmasq_loc=float(SW["hyp_sigmasq_loc"]), nu_s=float(SW["hyp_nu_s"]),
                                s_0=float(SW["hyp_s_0"]))
    got = np.array([lj[a] for a in ("Y", "x", "z", "v", "s")])
    assert np.all(np.abs(got - SW["r_logjoint"]) <= 1e-12 * np.abs(SW["r_logjoint"]))
    usage, durations = O.stateseq_stats(SW["z"], SW["mask"])
    assert np.array_equal(usage, SW["r_usage"]) and np.array_equal(durations, SW["r_durations"])


def _gpu_kw(SW):
    import torch
    t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a), device="cuda").to(dt)
    f32, f64 = torch.float32, torch.float64
    return dict(Y=t(SW["Y"], f32), mask=t(SW["mask"], f32), x=t(SW["x"], f32), v=t(SW["v"], f32), h=t(SW["h"], f32),
                s=t(SW["s"], f32), z=t(SW["z"], torch.int32), Ab=t(SW["Ab"], f64), Q=t(SW["Q"], f64),
                pi=t(SW["pi"], f64), Cd=t(SW["Cd"], f64), sigmasq=t(SW["sigmasq"], f64)), t


@both
@pytest.mark.gpu
def test_cuda_steps_equal_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    """The CUDA path against the reference's own outputs (same tolerances as tests/test_gpu_parity.py)."""
    import torch
    from keypointmoseq_b200 import gibbs as G
    kw, t = _gpu_kw(SW)
    f32, f64 = torch.float32, torch.float64
    n = lambda a: a.detach().cpu().numpy().astype(np.float64)
    # stateseqs: integer states bit-exact here (no near-ties in this fixture), log-likelihood 1e-5
    z, ll = G.resample_stateseqs(0, **kw, draws={"u": t(SW["d_stateseq_u"], f32)})
    valid = SW["mask"][:, L_:] > 0
    assert (n(z)[valid] == SW["r_stateseq_z"][valid]).mean() >= 0.999 and np.all(n(z)[~valid] == 0)
    assert np.all(np.abs(n(ll) - SW["r_stateseq_ll"]) <= 1e-5 * np.abs(SW["r_stateseq_ll"]))
    # latents
    x = G.resample_latents(0, **kw, draws={"w": t(SW["d_latents_w"], f32), "w_last": t(SW["d_latents_w_last"], f32)})
    xr = SW["r_latents_x"]
    assert np.abs(n(x) - xr).max() < 2e-6 * max(1.0, np.abs(xr).max())
    # heading
    # (in zero-padded frames the concentration is ~1e-14 and the mean direction is rounding noise)
    h = G.resample_heading(0, **kw, draws={"u": t(SW["d_heading_u"], f32)})
    live = SW["mask"] > 0
    assert np.abs(np.angle(np.exp(1j * (n(h) - SW["r_heading_h"]))))[live].max() < 1e-6
    # location
    v = G.resample_location(0, **kw, sigmasq_loc=float(SW["hyp_sigmasq_loc"]),
                            draws={"w": t(SW["d_location_w"], f32), "w_last": t(SW["d_location_w_last"], f32)})
    vr = SW["r_location_v"]
    assert np.abs(n(v) - vr).max() < 3e-7 * max(1.0, np.abs(vr).max())
    # scales
    s = G.resample_scales(0, **kw, nu_s=float(SW["hyp_nu_s"]), s_0=float(SW["hyp_s_0"]),
                          draws={"g": t(SW["d_scales_g"], f32)})
    assert np.allclose(n(s), SW["r_scales_s"], rtol=3e-7)
    # observation variance
    sg = G.resample_obs_variance(0, **kw, nu_sigma=float(SW["hyp_nu_sigma"]), sigmasq_0=float(SW["hyp_sigmasq_0"]),
                                 draws={"g": t(SW["d_obsvar_g"], f64)})
    assert np.allclose(n(sg), SW["r_obsvar_sigmasq"], rtol=1e-10)
    # AR parameters
    Ab, Q = G.resample_ar_params(0, nlags=L_, num_states=K_, mask=kw["mask"], x=kw["x"], z=kw["z"],
                                 nu_0=float(SW["hyp_nu_0"]), S_0=t(SW["hyp_S_0"], f64), M_0=t(SW["hyp_M_0"], f64),
                                 K_0=t(SW["hyp_K_0"], f64),
                                 draws={"chi2": t(SW["d_ar_chi2"], f64), "normals": t(SW["d_ar_normals"], f64),
                                        "G": t(SW["d_ar_G"], f64)})
    assert np.allclose(n(Q), SW["r_ar_Q"], rtol=1e-7, atol=1e-12)
    assert np.allclose(n(Ab), SW["r_ar_Ab"], rtol=1e-6, atol=1e-9)


@both
@pytest.mark.gpu
def test_cuda_obs_params_equals_reference_and_oracle(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    """SURVEY.md §8(f) rank 1: resample_obs_params (gibbs.py:58-82)."""
    import torch
    from keypointmoseq_b200 import gibbs as G
    kw, t = _gpu_kw(SW)
    Cd = G.resample_obs_params(0, Y=kw["Y"], mask=kw["mask"], sigmasq=kw["sigmasq"], v=kw["v"], h=kw["h"],
                               x=kw["x"], s=kw["s"], sigmasq_C=float(SW["hyp_sigmasq_C"]),
                               draws={"w": t(SW["d_obsparams_w"], torch.float64)})
    got = Cd.cpu().numpy()
    assert got.shape == SW["r_obsparams_Cd"].shape
    assert _rel(got, SW["r_obsparams_Cd"]) < 1e-9
    # the moments themselves, at a larger size, against the oracle
    from keypointmoseq_b200.synthetic import simulate
    d, s, p, Gm = simulate(9, 3, 500, k=16, dk=2, D=10, L=3, K=6, ragged=True, dtype=np.float32)
    f64 = lambda a: np.asarray(a, dtype=np.float64)
    Wo, Vo = O.obs_params_stats(Y=f64(d["Y"]), mask=f64(d["mask"]), sigmasq=p["sigmasq"], v=f64(s["v"]),
                                h=f64(s["h"]), x=f64(s["x"]), s=f64(s["s"]))
    c = lambda a: torch.as_tensor(a, device="cuda")
    W, V = G.obs_params_stats(Y=c(d["Y"]), mask=c(d["mask"]), sigmasq=c(p["sigmasq"]), v=c(s["v"]), h=c(s["h"]),
                              x=c(s["x"]), s=c(s["s"]))
    assert _rel(W.cpu().numpy(), Wo) < 1e-11 and _rel(V.cpu().numpy(), Vo) < 1e-11
    # default RNG path: finite, right shape
    Cd2 = G.resample_obs_params(3, Y=c(d["Y"]), mask=c(d["mask"]), sigmasq=c(p["sigmasq"]), v=c(s["v"]),
                                h=c(s["h"]), x=c(s["x"]), s=c(s["s"]), sigmasq_C=0.1)
    assert Cd2.shape == (15 * 2, 11) and torch.isfinite(Cd2).all()


@both
@pytest.mark.gpu
def test_cuda_diagnostics_equal_reference(fx):
    SW = FIX[fx]
    N_, T_, k_, dk_, D_, L_, K_, R_ = _dims(SW)
    """SURVEY.md §8(f) rank 2: log_joint_likelihood (util.py:232-275) and stateseq_stats (util.py:41-45)."""
    from keypointmoseq_b200 import gibbs as G
    kw, t = _gpu_kw(SW)
    lj = G.log_joint_likelihood(**kw, sigmasq_loc=float(SW["hyp_sigmasq_loc"]), nu_s=float(SW["hyp_nu_s"]),
                                s_0=float(SW["hyp_s_0"]))
    got = np.array([float(lj[a]) for a in ("Y", "x", "z", "v", "s")])
    assert np.all(np.abs(got - SW["r_logjoint"]) <= 1e-9 * np.abs(SW["r_logjoint"])), (got, SW["r_logjoint"])
    usage, durations = G.stateseq_stats(kw["z"], kw["mask"])
    assert np.array_equal(usage.cpu().numpy(), SW["r_usage"])
    assert np.array_equal(durations.cpu().numpy(), SW["r_durations"])


# ====================================================================== resample_hdp_transitions
# (bench fixture only) gibbs.py:148-151 -> transitions.py:23-40 run under the shim: numba's seeded
# CRP stream replayed as `randseq`, the two Dirichlet draws as normalised injected gammas; the
# reference's own sample_ms return value is recorded as r_hdp_ms.  That run showed that the binomial
# thinning of transitions.py:28-31 never reaches the returned counts (it is assigned to a temporary),
# so the default of oracle and product is the un-thinned table counts.
def _hdp_args(SW):
    return dict(alpha=float(SW["hyp_alpha"]), kappa=float(SW["hyp_kappa"]), gamma=float(SW["hyp_gamma"]))


def test_oracle_hdp_transitions_equals_reference():
    SW = FIX["bench"]
    K_ = _dims(SW)[6]
    betas, pi, counts, ms = O.resample_hdp_transitions(z=SW["z"], mask=SW["mask"], betas=SW["betas"], num_states=K_,
                                                       randseq=SW["d_hdp_randseq"],
                                                       betas_g=SW["d_hdp_betas_g"], pi_g=SW["d_hdp_pi_g"], **_hdp_args(SW))
    assert np.array_equal(counts, SW["r_hdp_counts"]) and np.array_equal(ms, SW["r_hdp_ms"])
    thinned = O.sample_ms(counts, SW["betas"], float(SW["hyp_alpha"]), float(SW["hyp_kappa"]), SW["d_hdp_randseq"],
                          SW["d_hdp_u_thin"], thin=True)
    assert (thinned.diagonal() < ms.diagonal()).any()      # the intended thinning would have changed them
    assert np.allclose(betas, SW["r_hdp_betas"], rtol=1e-14) and np.allclose(pi, SW["r_hdp_pi"], rtol=1e-14)
    # the injected gammas were drawn at the concentrations the reference then used
    a0, a1 = O.hdp_transition_concentrations(counts, ms, betas, float(SW["hyp_alpha"]), float(SW["hyp_kappa"]),
                                             float(SW["hyp_gamma"]))
    assert a0.shape == SW["d_hdp_betas_g"].shape and a1.shape == SW["d_hdp_pi_g"].shape and ms.sum() > 0


@pytest.mark.gpu
def test_cuda_hdp_transitions_equals_reference():
    import torch
    from keypointmoseq_b200 import gibbs as G
    SW = FIX["bench"]
    K_ = _dims(SW)[6]
    kw, t = _gpu_kw(SW)
    f32, f64 = torch.float32, torch.float64
    draws = {"randseq": t(SW["d_hdp_randseq"], f32), "betas_g": t(SW["d_hdp_betas_g"], f64),
             "pi_g": t(SW["d_hdp_pi_g"], f64)}
    # (the float32 cast of the uniforms flips no Bernoulli comparison in this fixture)
    betas, pi, info = G.resample_hdp_transitions(0, z=kw["z"], mask=kw["mask"], betas=t(SW["betas"], f64),
                                                 num_states=K_, draws=draws, return_info=True, **_hdp_args(SW))
    assert np.array_equal(info["ms"].cpu().numpy(), SW["r_hdp_ms"])
    assert np.allclose(betas.cpu().numpy(), SW["r_hdp_betas"], rtol=1e-13)
    assert np.allclose(pi.cpu().numpy(), SW["r_hdp_pi"], rtol=1e-13)
    # the Dirichlet concentrations the reference formed (transitions.py:38-39) from these statistics
    a0, a1 = O.hdp_transition_concentrations(SW["r_hdp_counts"], SW["r_hdp_ms"], SW["r_hdp_betas"],
                                             float(SW["hyp_alpha"]), float(SW["hyp_kappa"]), float(SW["hyp_gamma"]))
    assert np.allclose(info["betas_conc"].cpu().numpy(), a0, rtol=1e-14)
    assert np.allclose(info["pi_conc"].cpu().numpy(), a1, rtol=1e-13)
    counts = G.transition_counts(z=kw["z"], mask=kw["mask"], num_states=K_)
    assert np.array_equal(counts.cpu().numpy(), SW["r_hdp_counts"].astype(np.int64))
    # resample_transitions (gibbs.py:153-156) on the same counts
    conc = SW["r_hdp_counts"] + 5.0 + 30.0 * np.eye(K_)
    gp = np.random.default_rng(2).gamma(conc)
    pi2 = G.resample_transitions(0, z=kw["z"], mask=kw["mask"], alpha=5.0, kappa=30.0, num_states=K_,
                                 draws={"pi_g": t(gp, f64)})
    assert np.allclose(pi2.cpu().numpy(), O.sample_transitions(SW["r_hdp_counts"], 5.0, 30.0, gp)[0], rtol=1e-14)
    pi3 = G.resample_transitions(1, z=kw["z"], mask=kw["mask"], alpha=5.0, kappa=30.0, num_states=K_).cpu().numpy()
    assert np.allclose(pi3.sum(1), 1.0) and np.all(pi3 > 0)
