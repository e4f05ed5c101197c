"""Golden vectors from the reference's own gibbs.py / kalman.py / distributions.py / util.py,
executed in the build container under tests/golden/jax_numpy_shim.py (JAX itself is not
installed; see the shim's header for exactly what is and is not the reference's code).

Every resample_* step of keypoint_slds.ipynb cell 12 is run ONCE on the same problem, with injected
draws; inputs, draws and the reference's outputs go to
  reference_sweep.npz        small shapes (N=3, T=48, k=5, D=4, L=2, K=5), without resample_hdp_transitions
                             (its numba part is pinned separately by make_reference_pins.py);
  reference_sweep_bench.npz  the shapes the benchmark dispatches to (D=10, L=3 -> n=30 Kalman, 48-wide
                             emission tile; k=16; K=8; T=400, ragged), including resample_hdp_transitions
                             (numba's seeded stream replayed as the injected uniforms, the host-side
                             binomial thinning and the two Dirichlet draws injected, see the shim).

    python tests/golden/make_reference_sweep.py
"""
import importlib
import os
import sys

import numpy as np

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import jax_numpy_shim as shim                         # noqa: E402
from keypointmoseq_b200.synthetic import simulate, default_hypparams   # noqa: E402
from oracle import kpms_oracle as O                    # noqa: E402  (only its von Mises restatement)


def main(shape=(3, 48, 5, 2, 4, 2, 5, 6), fname="reference_sweep.npz", sim_seed=77, draw_seed=4242, with_hdp=False):
    shim.install()
    shim.VonMises.sampler = O.sample_vonmises
    G, U, TR = shim.import_reference("gibbs", "util", "transitions")

    N, T, k, dk, D, L, K, R = shape
    data, st, pr, Gamma = simulate(sim_seed, N, T, k=k, dk=dk, D=D, L=L, K=K, ragged=True, dtype=np.float64)
    trans, ar, obs, loc = default_hypparams(D, L, K)
    rng = np.random.default_rng(draw_seed)
    n = D * L
    f = lambda a: shim.sa(np.array(a, dtype=np.float64))
    r32 = lambda a: np.asarray(a, dtype=np.float32).astype(np.float64)   # float32-representable: the CUDA path stores these in fp32
    Y, mask, x, v, h, s = (f(r32(data["Y"])), f(data["mask"]), f(r32(st["x"])), f(r32(st["v"])), f(r32(st["h"])),
                           f(r32(st["s"])))
    z = shim.sa(np.array(st["z"], dtype=np.int64))
    Ab, Q, pi, Cd, sigmasq = f(pr["Ab"]), f(pr["Q"]), f(pr["pi"]), f(pr["Cd"]), f(pr["sigmasq"])
    key = shim.Key()
    out = dict(Y=Y, mask=mask, x=x, v=v, h=h, s=s, z=z, Ab=Ab, Q=Q, pi=pi, Cd=Cd, sigmasq=sigmasq,
               dims=np.array([N, T, k, dk, D, L, K, R]))
    for name, hp in (("ar", ar), ("obs", obs), ("loc", loc)):
        for a, b in hp.items():
            out[f"hyp_{a}"] = np.asarray(b, dtype=np.float64)

    def run(draws, fn, **kw):
        shim.Ctx.draws, shim.Ctx.vmap, shim.Ctx.scan_t = draws, [], None
        return fn(key, **kw)

    # 4: state sequences
    u = r32(rng.random((N, T - L)) * 0.998 + 0.001)
    zr, llr = run({"uniform": u}, G.resample_stateseqs, x=x, mask=mask, Ab=Ab, Q=Q, pi=pi)
    out.update(d_stateseq_u=u, r_stateseq_z=np.asarray(zr, dtype=np.int64), r_stateseq_ll=np.asarray(llr))
    # 5: latents
    w, wl = r32(rng.standard_normal((N, T - L, n))), r32(rng.standard_normal((N, n)))
    xr = run({"normal": w, "mvn": wl}, G.resample_latents, Y=Y, mask=mask, v=v, h=h, z=z, s=s, Cd=Cd,
             sigmasq=sigmasq, Ab=Ab, Q=Q)
    out.update(d_latents_w=w, d_latents_w_last=wl, r_latents_x=np.asarray(xr))
    # 6: heading
    uvm = r32(rng.random((N, T, R, 3)) * 0.998 + 0.001)
    hr = run({"vonmises": uvm}, G.resample_heading, Y=Y, v=v, x=x, s=s, Cd=Cd, sigmasq=sigmasq)
    out.update(d_heading_u=uvm, r_heading_h=np.asarray(hr))
    # 7: location
    wv, wvl = r32(rng.standard_normal((N, T - 1, dk))), r32(rng.standard_normal((N, dk)))
    vr = run({"normal": wv, "mvn": wvl}, G.resample_location, mask=mask, Y=Y, h=h, x=x, s=s, Cd=Cd,
             sigmasq=sigmasq, sigmasq_loc=float(loc["sigmasq_loc"]))
    out.update(d_location_w=wv, d_location_w_last=wvl, r_location_v=np.asarray(vr))
    # 8: scales
    gs = r32(rng.gamma((float(obs["nu_s"]) + 3) / 2, size=(N, T, k)))
    sr = run({"gamma": gs}, G.resample_scales, x=x, v=v, h=h, Y=Y, Cd=Cd, sigmasq=sigmasq,
             nu_s=float(obs["nu_s"]), s_0=float(obs["s_0"]))
    out.update(d_scales_g=gs, r_scales_s=np.asarray(sr))
    # 2: observation variance
    msum = float(mask.sum())
    gv = rng.gamma((float(obs["nu_sigma"]) + 3 * msum) / 2, size=(k,))
    sg = run({"gamma": gv}, G.resample_obs_variance, Y=Y, mask=mask, Cd=Cd, v=v, h=h, x=x, s=s,
             nu_sigma=float(obs["nu_sigma"]), sigmasq_0=float(obs["sigmasq_0"]))
    out.update(d_obsvar_g=gv, r_obsvar_sigmasq=np.asarray(sg))
    # 1: AR parameters (vmapped over states; chi-square draws enter as gamma(degs/2) * 2)
    counts = np.array([((np.asarray(z) == kk) & (np.asarray(mask)[:, L:] > 0)).sum() for kk in range(K)], dtype=np.float64)
    chi2 = np.stack([rng.chisquare(float(ar["nu_0"]) + counts[kk] - np.arange(D)) for kk in range(K)])
    nrm = rng.standard_normal((K, D * (D - 1) // 2))
    Gm = rng.standard_normal((K, D, D * L + 1))
    Abr, Qr = run({"gamma": chi2 / 2, "normal": nrm, "normal2": Gm}, G.resample_ar_params, nlags=L, num_states=K,
                  mask=mask, x=x, z=z, nu_0=float(ar["nu_0"]), S_0=f(ar["S_0"]), M_0=f(ar["M_0"]), K_0=f(ar["K_0"]))
    out.update(d_ar_chi2=chi2, d_ar_normals=nrm, d_ar_G=Gm, r_ar_Ab=np.asarray(Abr), r_ar_Q=np.asarray(Qr))

    # (f)1: observation matrix Cd (not part of the notebook's sweep; gibbs.py:58-82)
    wc = rng.standard_normal(dk * (D + 1) * (k - 1))
    Cdr = run({"mvn": wc}, G.resample_obs_params, Y=Y, mask=mask, sigmasq=sigmasq, v=v, h=h, x=x, s=s,
              sigmasq_C=float(obs["sigmasq_C"]))
    out.update(d_obsparams_w=wc, r_obsparams_Cd=np.asarray(Cdr))

    # (f)2: diagnostics (util.py:232-275, 41-45)
    shim.Ctx.draws, shim.Ctx.vmap, shim.Ctx.scan_t = {}, [], None
    lj = U.log_joint_likelihood(Y=Y, mask=mask, x=x, s=s, v=v, h=h, z=z, pi=pi, Ab=Ab, Q=Q, Cd=Cd, sigmasq=sigmasq,
                                sigmasq_loc=float(loc["sigmasq_loc"]), nu_s=float(obs["nu_s"]), s_0=float(obs["s_0"]))
    out["r_logjoint"] = np.array([float(lj[a]) for a in ("Y", "x", "z", "v", "s")])
    usage, durations = U.stateseq_stats(np.asarray(z), np.asarray(mask))
    out.update(r_usage=np.asarray(usage, dtype=np.int64), r_durations=np.asarray(durations, dtype=np.int64))

    if with_hdp:
        # 3: resample_hdp_transitions (gibbs.py:148-151 -> util.count_transitions (numba),
        # transitions.sample_hdp_transitions: numba CRP with its private seeded stream, host binomial
        # thinning, two jr.dirichlet draws)
        from numba import njit

        @njit
        def seed_numba(sd):
            np.random.seed(sd)

        # kappa chosen so that the binomial thinning of the diagonal keeps a real fraction of the tables
        alpha, kappa, gamma = float(trans["alpha"]), 50.0, float(trans["gamma"])
        betas = f(pr["betas"])
        crp_seed = 9091
        counts_ref = np.asarray(U.count_transitions(K, np.array(mask[..., -z.shape[-1]:]), np.array(z)))
        tot = int(counts_ref.sum())
        randseq = np.random.RandomState(crp_seed).random_sample(tot)       # = numba's stream after seed_numba(crp_seed)
        u_thin = rng.random((K, 4096))
        rho_all = np.asarray(betas) * alpha / (np.asarray(betas) * alpha + kappa)
        # what the reference's sample_ms returns: the raw CRP counts (its thinning writes to a temporary)
        ms_expected = O.sample_ms(counts_ref.astype(np.int64), np.asarray(betas), alpha, kappa, randseq)
        gb = rng.gamma(ms_expected.sum(0) + gamma / K)
        gp = rng.gamma(alpha * (gb / gb.sum())[None, :] + counts_ref + kappa * np.eye(K))
        real_binomial, real_sample_ms, seen = np.random.binomial, TR.sample_ms, {}

        def recording_sample_ms(*a, **kw):       # the reference's own sample_ms; its return value is recorded
            seen["ms"] = np.asarray(real_sample_ms(*a, **kw)).copy()
            return seen["ms"]
        TR.sample_ms = recording_sample_ms
        np.random.binomial = shim.bernoulli_sum_binomial(u_thin, rho_all)
        try:
            seed_numba(crp_seed)
            shim.Ctx.draws, shim.Ctx.vmap, shim.Ctx.scan_t = {"dirichlet_g": [gb, gp]}, [], None
            betas_r, pi_r = G.resample_hdp_transitions(key, z=z, mask=mask, betas=betas, alpha=alpha, kappa=kappa,
                                                       gamma=gamma, num_states=K)
        finally:
            np.random.binomial, TR.sample_ms = real_binomial, real_sample_ms
        assert np.array_equal(seen["ms"], ms_expected), "the injected Dirichlet gammas were drawn for other table counts"
        out.update(hyp_alpha=np.float64(alpha), hyp_kappa=np.float64(kappa), hyp_gamma=np.float64(gamma),
                   betas=np.asarray(betas), d_hdp_randseq=randseq, d_hdp_u_thin=u_thin, d_hdp_betas_g=gb,
                   d_hdp_pi_g=gp, r_hdp_counts=counts_ref, r_hdp_ms=seen["ms"], r_hdp_betas=np.asarray(betas_r), r_hdp_pi=np.asarray(pi_r))

    path = os.path.join(HERE, fname)
    np.savez_compressed(path, **{a: np.asarray(b) for a, b in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
    main(shape=(3, 400, 16, 2, 10, 3, 8, 8), fname="reference_sweep_bench.npz", sim_seed=78, draw_seed=4343,
         with_hdp=True)
