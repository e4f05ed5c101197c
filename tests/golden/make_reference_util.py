"""Golden vectors for the helpers around the sweep: the reference's own util.py (layout helpers,
merge_data, interpolate_keypoints) and initialize.py (initial_*), executed in the build container
under tests/golden/jax_numpy_shim.py (see its header for what is the reference's code and what is
injected).  sklearn's PCA, which initialize.py:25 calls, is installed here and runs as it is.

    python tests/golden/make_reference_util.py     # writes tests/golden/reference_util.npz
"""
import os
import sys

import numpy as np

sys.dont_write_bytecode = True
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import jax_numpy_shim as shim                                            # noqa: E402
from keypointmoseq_b200.synthetic import simulate, default_hypparams     # noqa: E402


def main():
    shim.install()
    U, I = shim.import_reference("util", "initialize")
    rng = np.random.default_rng(515)
    f = lambda a: shim.sa(np.array(a, dtype=np.float64))
    out = {}

    # ---- util.py layout helpers
    x = f(rng.standard_normal((2, 9, 3)))
    out.update(u_x=x, u_get_lags=U.get_lags(x, 2), u_pad_affine=U.pad_affine(x))
    K, D, L, p = 3, 4, 3, 6
    As, bs, Qs, Cs = f(rng.standard_normal((K, D, D * L))), f(rng.standard_normal((K, D))), \
        f(rng.standard_normal((K, D, D))), f(rng.standard_normal((p, D)))
    A_, b_, Q_, C_ = U.ar_to_lds(As, bs, Qs, Cs)
    out.update(u_As=As, u_bs=bs, u_Qs=Qs, u_Cs=Cs, u_lds_A=A_, u_lds_b=b_, u_lds_Q=Q_, u_lds_C=C_)
    V = f(rng.standard_normal((5, 7, 2)))
    V[0, 0] = (-1.0, 0.0)      # the branch cut
    V[0, 1] = (0.0, 1.0)
    hh = f(rng.uniform(-np.pi, np.pi, size=(2, 5)))
    out.update(u_V=V, u_vector_to_angle=U.vector_to_angle(V), u_h=hh,
               u_rot2=U.angle_to_rotation_matrix(hh, keypoint_dim=2), u_rot3=U.angle_to_rotation_matrix(hh))
    for d in (2, 3):
        Yd, vd = f(rng.standard_normal((2, 5, 4, d))), f(rng.standard_normal((2, 5, d)))
        out.update({f"u_Y{d}": Yd, f"u_v{d}": vd, f"u_affine{d}": U.affine_transform(Yd, vd, hh),
                    f"u_inv_affine{d}": U.inverse_affine_transform(Yd, vd, hh)})
    k = 4
    Gam = np.asarray(U.center_embedding(k))
    xx, Cd = f(rng.standard_normal((2, 5, 3))), f(rng.standard_normal(((k - 1) * 2, 4)))
    out.update(u_Gamma=Gam, u_ek_x=xx, u_ek_Cd=Cd,
               u_expected_keypoints=U.expected_keypoints(Y=out["u_Y2"], v=out["u_v2"], h=hh, x=xx, Cd=Cd))
    comps, means = f(rng.standard_normal((8, 3))), f(rng.standard_normal(8))
    out.update(u_pca_components=comps, u_pca_means=means,
               u_pca_transform=U.pca_transform(comps, means, Y=out["u_Y2"], v=out["u_v2"], h=hh))
    X = f(rng.standard_normal((3, 4, 4)))
    out.update(u_X=X, u_ensure_symmetric=U.ensure_symmetric(X))

    # ---- ingest: merge_data (ragged sessions), interpolate_keypoints
    lens = {"b_sess": 37, "a_sess": 50, "c_sess": 12}
    dd = {nm: rng.standard_normal((t, 4, 2)) for nm, t in lens.items()}
    for nm in sorted(lens):
        out["m_in_" + nm] = dd[nm]
    for tag, bl in (("full", None), ("cut", 20)):
        data, mask, keys = U.merge_data(dd, batch_length=bl)
        out.update({f"m_{tag}_data": data, f"m_{tag}_mask": mask,
                    f"m_{tag}_keys": np.array([f"{a}:{b}" for a, b in keys])})
    kp = rng.standard_normal((3, 40, 5, 2))
    outl = rng.random((3, 40, 5)) < 0.25
    outl[0, :4, 0] = True      # leading run
    outl[1, -3:, 2] = True     # trailing run
    out.update(i_kp=kp.copy(), i_outliers=outl, i_interp=U.interpolate_keypoints(kp.copy(), outl, axis=1))

    # ---- initialize.py
    n, t, kk, d3 = 3, 60, 6, 3
    Y3 = f(rng.standard_normal((n, t, kk, d3)) + rng.standard_normal((n, 1, 1, d3)) * 3)
    post, ante = np.array([0, 1]), np.array([4, 5])
    out.update(n_Y3=Y3, n_post=post, n_ante=ante, n_location=I.initial_location(Y=Y3),
               n_heading=I.initial_heading(post, ante, Y=Y3))
    N, T, k2, D2 = 3, 300, 8, 5
    data, st, pr, Gamma = simulate(5, N, T, k=k2, dk=2, D=D2, L=3, K=4, ragged=True, dtype=np.float64)
    Y2, mask2, v2, h2 = f(data["Y"]), f(data["mask"]), f(st["v"]), f(st["h"])
    seed, ns = 1234, 500
    np.random.seed(seed)
    lat, Cd0, pca = I.initial_latents(Y=Y2, mask=mask2, v=v2, h=h2, latent_dim=D2, num_samples=ns)
    out.update(n_Y2=Y2, n_mask2=mask2, n_v2=v2, n_h2=h2, n_Gamma=np.asarray(U.center_embedding(k2)),
               n_sample_index=np.random.RandomState(seed).choice(T * N, ns), n_latents=lat, n_Cd=Cd0,
               n_pca_components=pca.components_, n_pca_mean=pca.mean_,
               n_pca_evr=pca.explained_variance_ratio_)
    np.random.seed(seed)
    lat_nw, Cd_nw, _ = I.initial_latents(Y=Y2, mask=mask2, v=v2, h=h2, latent_dim=D2, num_samples=ns, whiten=False)
    out.update(n_latents_nowhiten=lat_nw, n_Cd_nowhiten=Cd_nw)

    Kst, Dd, Ll = 6, 4, 2
    trans, ar, obs, loc = default_hypparams(Dd, Ll, Kst)
    key = shim.Key()
    chi2 = np.stack([rng.chisquare(float(ar["nu_0"]) - np.arange(Dd)) for _ in range(Kst)])
    nrm, Gm = rng.standard_normal((Kst, Dd * (Dd - 1) // 2)), rng.standard_normal((Kst, Dd, Dd * Ll + 1))
    shim.Ctx.draws, shim.Ctx.vmap, shim.Ctx.scan_t = {"gamma": chi2 / 2, "normal": nrm, "normal2": Gm}, [], None
    Ab0, Q0 = I.initial_ar_params(key, num_states=Kst, nu_0=float(ar["nu_0"]), S_0=f(ar["S_0"]), M_0=f(ar["M_0"]),
                                  K_0=f(ar["K_0"]))
    out.update(n_ar_nu_0=np.float64(ar["nu_0"]), n_ar_S_0=ar["S_0"], n_ar_M_0=ar["M_0"], n_ar_K_0=ar["K_0"],
               n_ar_chi2=chi2, n_ar_normals=nrm, n_ar_G=Gm, n_ar_Ab=Ab0, n_ar_Q=Q0)
    alpha, kappa, gamma = 5.7, 1e3, 1e3
    g0 = rng.gamma(np.ones(Kst) * gamma / Kst)
    g1 = rng.gamma(np.zeros(Kst) + gamma / Kst)                      # ms = 0 on zero counts
    b1 = g1 / g1.sum()
    g2 = rng.gamma(alpha * b1[None, :] + kappa * np.eye(Kst))
    shim.Ctx.draws = {"dirichlet_g": [g0, g1, g2]}
    betas0, pi0 = I.initial_hdp_transitions(key, num_states=Kst, alpha=alpha, kappa=kappa, gamma=gamma)
    g3 = rng.gamma(alpha + kappa * np.eye(Kst))
    shim.Ctx.draws = {"dirichlet_g": [g3]}
    pi1 = I.initial_transitions(key, num_states=Kst, alpha=alpha, kappa=kappa)
    out.update(n_hdp=np.array([alpha, kappa, gamma]), n_hdp_g0=g0, n_hdp_g1=g1, n_hdp_g2=g2, n_hdp_betas=betas0,
               n_hdp_pi=pi0, n_trans_g=g3, n_trans_pi=pi1)

    path = os.path.join(HERE, "reference_util.npz")
    np.savez_compressed(path, **{a: np.asarray(b) for a, b in out.items()})
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
