"""Generate tests/golden/sweep_small.npz: seeded inputs, injected draws and the oracle's
outputs for every hot-path row at a tiny size.  The reference itself cannot run here (no JAX),
so the vectors come from the oracle restatement; they freeze its behaviour so that later
edits to oracle or kernels are caught.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import kpms_helpers as H  # noqa: E402
from oracle import kpms_oracle as O  # noqa: E402


def main():
    N, T, k, dk, D, L, K = 2, 48, 5, 2, 10, 3, 4
    d, s, p, Gm = H.problem(seed=77, N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, ragged=True)
    rng = np.random.default_rng(78)
    dr = H.sweep_draws(rng, N, T, k, dk, D, L, K, rounds=6)
    hp = H.hyp(D, L, K)
    out = dict(N=N, T=T, k=k, dk=dk, D=D, L=L, K=K, Gamma=Gm)
    out.update({"data_" + a: b for a, b in d.items()})
    out.update({"st_" + a: b for a, b in s.items()})
    out.update({"par_" + a: b for a, b in p.items()})
    out.update({"dr_" + a: b for a, b in dr.items()})
    out["ll"] = O.ar_log_likelihood(s['x'], p['Ab'], p['Q'])
    out["z"], out["z_loglik"] = O.resample_stateseqs(x=s['x'], mask=d['mask'], Ab=p['Ab'], Q=p['Q'], pi=p['pi'], u=dr['z_u'])
    out["x"], out["x_loglik"] = O.resample_latents(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], z=s['z'], s=s['s'],
                                                   Cd=p['Cd'], sigmasq=p['sigmasq'], Ab=p['Ab'], Q=p['Q'], Gamma=Gm,
                                                   w=dr['x_w'], w_last=dr['x_w_last'], return_loglik=True)
    out["F"] = O.ar_sufficient_stats(nlags=L, num_states=K, mask=d['mask'], x=s['x'], z=s['z'])
    out["counts"] = O.count_transitions(K, d['mask'][:, -s['z'].shape[1]:], s['z'])
    oc = dict(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], x=s['x'], s=s['s'], Cd=p['Cd'], sigmasq=p['sigmasq'], Gamma=Gm)
    out["S_y"], out["mask_sum"] = O.obs_variance_stats(**oc)
    out["theta"], out["kappa"] = O.heading_stats(**oc)
    out["h"] = O.sample_vonmises(out["theta"], out["kappa"], dr['h_u'])
    out["v"], out["v_loglik"] = O.resample_location(**oc, sigmasq_loc=0.5, w=dr['v_w'], w_last=dr['v_w_last'], return_loglik=True)
    g = H.f32r(rng.gamma((hp['nu_s'] + 3) / 2, size=(N, T, k)))
    out["dr_s_g"] = g
    out["s_new"] = O.resample_scales(**oc, nu_s=hp['nu_s'], s_0=hp['s_0'], g=g)
    np.savez_compressed(os.path.join(HERE, "sweep_small.npz"), **out)
    print("wrote", os.path.join(HERE, "sweep_small.npz"))


if __name__ == "__main__":
    main()
