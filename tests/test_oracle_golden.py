"""The oracle must keep reproducing the committed golden vectors (tests/golden/), and the CUDA
path must match them too (GPU-marked test)."""
import os

import numpy as np
import pytest

import kpms_helpers as H
from oracle import kpms_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sweep_small.npz")


@pytest.fixture(scope="module")
def g():
    return dict(np.load(GOLD))


def _oc(g):
    return dict(Y=g['data_Y'], mask=g['data_mask'], v=g['st_v'], h=g['st_h'], x=g['st_x'], s=g['st_s'],
                Cd=g['par_Cd'], sigmasq=g['par_sigmasq'], Gamma=g['Gamma'])


def test_oracle_reproduces_golden(g):
    L, K = int(g['L']), int(g['K'])
    assert np.allclose(O.ar_log_likelihood(g['st_x'], g['par_Ab'], g['par_Q']), g['ll'], rtol=1e-12, atol=1e-12)
    z, ll = O.resample_stateseqs(x=g['st_x'], mask=g['data_mask'], Ab=g['par_Ab'], Q=g['par_Q'], pi=g['par_pi'],
                                 u=g['dr_z_u'])
    assert np.array_equal(z, g['z']) and np.allclose(ll, g['z_loglik'], rtol=1e-12)
    x, lx = O.resample_latents(Y=g['data_Y'], mask=g['data_mask'], v=g['st_v'], h=g['st_h'], z=g['st_z'], s=g['st_s'],
                               Cd=g['par_Cd'], sigmasq=g['par_sigmasq'], Ab=g['par_Ab'], Q=g['par_Q'],
                               Gamma=g['Gamma'], w=g['dr_x_w'], w_last=g['dr_x_w_last'], return_loglik=True)
    assert np.allclose(x, g['x'], rtol=1e-9, atol=1e-10) and np.allclose(lx, g['x_loglik'], rtol=1e-11)
    F = O.ar_sufficient_stats(nlags=L, num_states=K, mask=g['data_mask'], x=g['st_x'], z=g['st_z'])
    assert np.allclose(F, g['F'], rtol=1e-12, atol=1e-12)
    assert np.array_equal(O.count_transitions(K, g['data_mask'][:, -g['st_z'].shape[1]:], g['st_z']), g['counts'])
    v, lv = O.resample_location(**_oc(g), sigmasq_loc=0.5, w=g['dr_v_w'], w_last=g['dr_v_w_last'], return_loglik=True)
    assert np.allclose(v, g['v'], rtol=1e-10, atol=1e-10) and np.allclose(lv, g['v_loglik'], rtol=1e-11)


@pytest.mark.gpu
def test_cuda_path_matches_golden(g):
    import torch
    from keypointmoseq_b200 import gibbs as G
    T_ = lambda a: torch.as_tensor(np.ascontiguousarray(a), device="cuda")
    N_ = lambda t: t.detach().cpu().numpy()
    L, K = int(g['L']), int(g['K'])
    z, ll = G.resample_stateseqs(0, x=T_(g['st_x']), mask=T_(g['data_mask']), Ab=T_(g['par_Ab']), Q=T_(g['par_Q']),
                                 pi=T_(g['par_pi']), draws={'u': T_(g['dr_z_u'])})
    valid = g['data_mask'][:, L:] > 0
    assert (N_(z)[valid] == g['z'][valid]).mean() >= 0.999
    assert np.all(np.abs(N_(ll) - g['z_loglik']) <= 1e-5 * np.abs(g['z_loglik']))
    x, info = G.resample_latents(0, Y=T_(g['data_Y']), mask=T_(g['data_mask']), v=T_(g['st_v']), h=T_(g['st_h']),
                                 z=T_(g['st_z'].astype(np.int32)), s=T_(g['st_s']), Cd=T_(g['par_Cd']),
                                 sigmasq=T_(g['par_sigmasq']), Ab=T_(g['par_Ab']), Q=T_(g['par_Q']), Gamma=T_(g['Gamma']),
                                 draws={'w': T_(g['dr_x_w']), 'w_last': T_(g['dr_x_w_last'])}, return_info=True)
    assert np.abs(N_(x) - g['x']).max() < 2e-6 * max(1.0, np.abs(g['x']).max())
    assert np.all(np.abs(N_(info['loglik']) - g['x_loglik']) <= 1e-9 * np.abs(g['x_loglik']))
    F = N_(G.ar_sufficient_stats(nlags=L, num_states=K, mask=T_(g['data_mask']), x=T_(g['st_x']),
                                 z=T_(g['st_z'].astype(np.int32))))
    assert np.abs(F - g['F']).max() <= 1e-12 * np.abs(g['F']).max()
    c = N_(G.transition_counts(z=T_(g['st_z'].astype(np.int32)), mask=T_(g['data_mask']), num_states=K))
    assert np.array_equal(c, g['counts'].astype(np.int64))
    common = dict(Y=T_(g['data_Y']), mask=T_(g['data_mask']), v=T_(g['st_v']), h=T_(g['st_h']), x=T_(g['st_x']),
                  s=T_(g['st_s']), Cd=T_(g['par_Cd']), sigmasq=T_(g['par_sigmasq']), Gamma=T_(g['Gamma']))
    S, m = G.obs_variance_stats(**common)
    assert np.allclose(N_(S), g['S_y'], rtol=1e-11) and float(m) == float(g['mask_sum'])
    hh = G.resample_heading(0, **common, draws={'u': T_(g['dr_h_u'])})
    ok = g['kappa'] > 1e-6   # zero-padded frames have kappa ~ 1e-14: the mean direction is rounding noise there
    assert np.abs(np.angle(np.exp(1j * (N_(hh) - g['h']))))[ok].max() < 1e-6
    v = G.resample_location(0, **common, sigmasq_loc=0.5, draws={'w': T_(g['dr_v_w']), 'w_last': T_(g['dr_v_w_last'])})
    assert np.abs(N_(v) - g['v']).max() < 3e-7 * max(1.0, np.abs(g['v']).max())
    hp = H.hyp(int(g['D']), L, K)
    sn = G.resample_scales(0, **common, nu_s=hp['nu_s'], s_0=hp['s_0'], draws={'g': T_(g['dr_s_g'])})
    assert np.allclose(N_(sn), g['s_new'], rtol=3e-7)
