"""The associative-scan variant of the Kalman filter / sampler (`keypointmoseq_b200/kalman_scan.py`,
BASELINE north_star (3)): exact against the oracle on the CPU, and the exact reference point of the
chunk-parallel CUDA kernels on the GPU (same inputs, same injected draws)."""
import numpy as np
import pytest
import torch

import kpms_helpers as H
from oracle import kpms_oracle as O
from keypointmoseq_b200 import kalman_scan as KS


def _case(seed, N, T, k, D, L, K, ragged=True):
    d, s, p, Gm = H.problem(seed=seed, N=N, T=T, k=k, dk=2, D=D, L=L, K=K, ragged=ragged)
    rng = np.random.default_rng(seed + 1)
    w = H.f32r(rng.standard_normal((N, T - L, D * L)))
    wl = H.f32r(rng.standard_normal((N, D * L)))
    return d, s, p, Gm, w, wl


def _scan(d, s, p, Gm, w, wl, dev):
    t = lambda a: torch.as_tensor(a, device=dev)
    return KS.resample_latents_scan(0, Y=t(d['Y']), mask=t(d['mask']), v=t(s['v']), h=t(s['h']), z=t(s['z']),
                                    s=t(s['s']), Cd=t(p['Cd']), sigmasq=t(p['sigmasq']), Ab=t(p['Ab']), Q=t(p['Q']),
                                    Gamma=t(Gm), draws={'w': t(w), 'w_last': t(wl)}, return_info=True)


def test_associative_scan_is_a_prefix_product():
    """2 x 2 matrix products (non-commutative), every length from 1 to 19, forwards and backwards."""
    rng = np.random.default_rng(0)
    mul = lambda a, b: (a[0] @ b[0],)
    for T in range(1, 20):
        M = torch.as_tensor(rng.standard_normal((T, 3, 2, 2)))
        (fw,) = KS.associative_scan(mul, (M,))
        (bw,) = KS.associative_scan(lambda a, b: (b[0] @ a[0],), (M,), reverse=True)
        acc = M[0]
        for t in range(T):
            acc = M[t] if t == 0 else acc @ M[t]
            assert torch.allclose(fw[t], acc, atol=1e-12)
        acc = M[T - 1]
        for t in range(T - 1, -1, -1):
            acc = M[t] if t == T - 1 else M[t] @ acc
            assert torch.allclose(bw[t], acc, atol=1e-12)


@pytest.mark.parametrize("N,T,k,D,L,K", [(2, 60, 5, 4, 2, 3), (3, 257, 8, 10, 3, 6)])
def test_scan_variant_equals_oracle(N, T, k, D, L, K):
    d, s, p, Gm, w, wl = _case(5, N, T, k, D, L, K)
    xo, llo = O.resample_latents(Y=d['Y'], mask=d['mask'], v=s['v'], h=s['h'], z=s['z'], s=s['s'], Cd=p['Cd'],
                                 sigmasq=p['sigmasq'], Ab=p['Ab'], Q=p['Q'], Gamma=Gm, w=w, w_last=wl,
                                 return_loglik=True)
    x, info = _scan(d, s, p, Gm, w, wl, "cpu")
    assert np.abs(x.numpy() - xo).max() < 1e-10 * max(1.0, np.abs(xo).max())
    assert np.all(np.abs(info['loglik'].numpy() - llo) <= 1e-12 * np.abs(llo))


@pytest.mark.gpu
@pytest.mark.parametrize("chunk_len,burn_in", [(128, 160), ("auto", 256), (0, 0)])
def test_chunked_cuda_kernels_against_the_scan(chunk_len, burn_in):
    """The chunk-with-warm-up kernels and the exact scan see the same data and draws (n = 30)."""
    from keypointmoseq_b200 import gibbs as G
    N, T, k, D, L, K = 2, 1500, 16, 10, 3, 8
    d, s, p, Gm, w, wl = _case(9, N, T, k, D, L, K)
    xs, info_s = _scan(d, s, p, Gm, w, wl, "cuda")
    t = lambda a: torch.as_tensor(a, device="cuda")
    x, info = G.resample_latents(0, Y=t(d['Y']), mask=t(d['mask']), v=t(s['v']), h=t(s['h']), z=t(s['z']),
                                 s=t(s['s']), Cd=t(p['Cd']), sigmasq=t(p['sigmasq']), Ab=t(p['Ab']), Q=t(p['Q']),
                                 Gamma=t(Gm), draws={'w': t(w), 'w_last': t(wl)}, chunk_len=chunk_len,
                                 burn_in=burn_in, return_info=True)
    scale = max(1.0, float(xs.abs().max()))
    assert float((x.double() - xs).abs().max()) < 2e-6 * scale
    ls, lc = info_s['loglik'].cpu().numpy(), info['loglik'].cpu().numpy()
    assert np.all(np.abs(ls - lc) <= 1e-9 * np.abs(ls)), (ls, lc)


def test_generic_kalman_entry_points_equal_oracle():
    """`keypoint_moseq.kalman.kalman_filter / kalman_sample` (one sequence, generic LDS) against the
    oracle's restatement of kalman.py:80-156, with a masked stretch in the middle of the sequence."""
    from keypoint_moseq.kalman import kalman_filter, kalman_sample
    rng = np.random.default_rng(3)
    Tp, n, p, K = 40, 5, 4, 3
    A = 0.8 * np.eye(n)[None] + 0.1 * rng.standard_normal((K, n, n))
    B = 0.1 * rng.standard_normal((K, n))
    Q = np.stack([np.eye(n) * 0.3 + 0.05 * (lambda m: m @ m.T)(rng.standard_normal((n, n))) for _ in range(K)])
    C, D = rng.standard_normal((p, n)), rng.standard_normal(p)
    Rs = 0.5 + rng.random((Tp, p))
    ys = rng.standard_normal((Tp, p))
    zs = rng.integers(0, K, Tp - 1)
    mask = np.ones(Tp - 1)
    mask[17:23] = 0
    m0, S0 = np.zeros(n), 10 * np.eye(n)
    w, wl = rng.standard_normal((Tp - 1, n)), rng.standard_normal(n)
    ms_o, Ss_o, _ = O.kalman_filter(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs)
    xs_o, _ = O.kalman_sample(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, w, wl)
    ms, Ss = kalman_filter(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs)
    xs = kalman_sample(0, ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, draws={'w': w, 'w_last': wl})
    assert np.abs(ms.numpy() - ms_o).max() < 1e-10 and np.abs(Ss.numpy() - Ss_o).max() < 1e-10
    assert np.abs(xs.numpy() - xs_o).max() < 1e-10
