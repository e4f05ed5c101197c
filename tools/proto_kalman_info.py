"""Design prototype (NumPy): information-form Kalman filter / backward sampler used
by the CUDA kernel, checked against the covariance-form oracle (kalman.py restated).
Run: python tools/proto_kalman_info.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import kpms_oracle as O


def rev_chol_inv(Lam):
    """L lower with L L^T = inv(Lam): Lam = M^T M (M lower, computed bottom-up), L = M^-1."""
    n = Lam.shape[0]
    W = Lam.copy()
    M = np.zeros_like(W)
    for k in range(n - 1, -1, -1):
        M[k, k] = np.sqrt(W[k, k])
        M[k, :k] = W[k, :k] / M[k, k]
        W[:k, :k] -= np.outer(M[k, :k], M[k, :k])
    return np.linalg.inv(M)


def info_kalman_sample(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, w, w_last):
    Tp, p = ys.shape
    n = m0.shape[0]
    K = A.shape[0]
    Qinv = np.linalg.inv(Q)
    F = np.einsum('kji,kjl->kil', A, Qinv)          # A^T Qinv
    G = F @ A
    f = np.einsum('kij,kj->ki', F, B)
    qb = np.einsum('kij,kj->ki', Qinv, B)
    cz = np.einsum('ki,ki->k', B, qb)
    ldq = np.linalg.slogdet(Q)[1]
    Lam = np.linalg.inv(S0); eta = Lam @ m0
    ll = 0.5 * np.linalg.slogdet(Lam)[1] - 0.5 * m0 @ Lam @ m0
    Ls = np.zeros((Tp, n, n)); etas = np.zeros((Tp, n))
    for t in range(Tp - 1):
        if mask[t] > 0:
            yt = ys[t] - D
            CtR = C.T / Rs[t]
            ll += -0.5 * np.log(2 * np.pi * Rs[t]).sum() - 0.5 * (yt * yt / Rs[t]).sum()
            Lc = Lam + CtR @ C; ec = eta + CtR @ yt
            z = zs[t]
            L = rev_chol_inv(Lc + G[z])
            Ls[t] = L; etas[t] = ec
            u = L.T @ (ec - f[z]); V = L.T @ F[z]
            Lam = Qinv[z] - V.T @ V; eta = qb[z] + V.T @ u
            ll += -0.5 * ldq[z] + np.log(np.diag(L)).sum() + 0.5 * (u @ u - cz[z])
    yt = ys[-1] - D; CtR = C.T / Rs[-1]
    ll += -0.5 * np.log(2 * np.pi * Rs[-1]).sum() - 0.5 * (yt * yt / Rs[-1]).sum()
    Lc = Lam + CtR @ C; ec = eta + CtR @ yt
    L = rev_chol_inv(Lc)
    tl = L.T @ ec
    ll += np.log(np.diag(L)).sum() + 0.5 * tl @ tl
    xs = np.zeros((Tp, n))
    x = L @ (tl + w_last); xs[-1] = x
    for t in range(Tp - 2, -1, -1):
        if mask[t] > 0:
            z = zs[t]
            h = etas[t] + F[z] @ (x - B[z])
            x = Ls[t] @ (Ls[t].T @ h + w[t]); xs[t] = x
    return xs, ll


if __name__ == '__main__':
    from keypointmoseq_b200.synthetic import simulate
    data, st, pr, Gamma = simulate(0, 2, 400, k=8, dk=2, D=10, L=3, K=6, ragged=True, dtype=np.float64)
    Y, mask = data['Y'], data['mask']
    d, L = 10, 3
    Cdf = np.kron(Gamma, np.eye(2)) @ pr['Cd']
    ys = O.inverse_affine_transform(Y, st['v'], st['h']).reshape(2, 400, -1)
    A, B, Q_, C = O.ar_to_lds(pr['Ab'][..., :-1], pr['Ab'][..., -1], pr['Q'], Cdf[..., :-1])
    Dv = Cdf[..., -1]
    R = np.repeat(st['s'] * pr['sigmasq'], 2, axis=-1)[:, L - 1:]
    rng = np.random.default_rng(1)
    for i in range(2):
        w = rng.standard_normal((400 - L, 30)); wl = rng.standard_normal(30)
        a = (ys[i, L - 1:], mask[i, L - 1:-1], st['z'][i], np.zeros(30), np.eye(30) * 10, A, B, Q_, C, Dv, R[i], w, wl)
        xo, llo = O.kalman_sample(*a)
        xi, lli = info_kalman_sample(*a)
        print('seq', i, 'max|dx|', np.abs(xo - xi).max(), 'll', llo, lli, 'rel', abs(llo - lli) / abs(llo))
