"""Exact time-parallel variant of the Kalman filter / backward sampler: an ASSOCIATIVE SCAN over time.

`BASELINE.json` north_star (3) names "an associative parallel-scan variant over time" next to the
chunk-per-CTA kernels.  The hand-written kernels (`csrc/kalman.cu`) cut a session into chunks with a
verified warm-up; this module is the other variant: no chunks, no warm-up, no tolerance — prefix
products of the elements of temporal parallelisation of Bayesian filters (Sarkka & Garcia-Fernandez,
IEEE TAC 2021), depth log2(T).  It exists as the exact reference point that cross-checks the chunked
kernels at the same inputs and draws (`tests/test_kalman_scan.py`), and for callers who want the
time-parallel form.  It is NOT on the product path of `gibbs.resample_latents` and is never chosen
implicitly: every combine is a batch of dense n x n float64 solves and products handed to
cuBLAS / cuSOLVER through torch (library kernels, like the MNIW draws), 2-3x the flops of a
sequential step per level and 22 KB per element, which is why the n = 30 filter of the sweep stays
chunked.

Reference semantics reproduced (kalman.py:80-156): the carry
is the PREDICTED moment; step t conditions on y_t and then predicts with z_t; masked steps leave
the carry untouched and emit zeros; the final update is unconditional.

Element of step t (unmasked), acting on the predicted state x_t:
    p(x_{t+1} | x_t)        = N(A x_t + b, C)          A = A_z, b = B_z, C = Q_z
    p(y_t | x_t)  propto  exp(-1/2 x^T J x + eta^T x)  J = C_y^T R_t^-1 C_y, eta = C_y^T R_t^-1 (y_t - D)
masked step: (I, 0, 0, 0, 0); the prior is the element (0, m0, S0, 0, 0) in front.  Combination
(i earlier, j later), with M = (I + C_i J_j)^-1:
    A = A_j M A_i                  b = A_j M (b_i + C_i eta_j) + b_j     C = A_j M C_i A_j^T + C_j
    eta = A_i^T M^T (eta_j - J_j b_i) + eta_i                            J = A_i^T M^T J_j A_i + J_i
The backward sampler is a reverse scan of affine maps x_t = G_t x_{t+1} + c_t.
"""
import math

import torch

from . import util


def associative_scan(combine, elems, reverse=False):
    """Inclusive scan along dim 0 of a tuple of tensors; `combine(a, b)` with `a` the EARLIER
    operand (for reverse=True: `a` the later one, i.e. the scan runs from the end).  Recursive
    pairing (work-efficient, depth 2 log2 T), every level one batched call of `combine`."""
    if reverse:
        out = associative_scan(combine, tuple(e.flip(0) for e in elems))
        return tuple(e.flip(0) for e in out)
    T = elems[0].shape[0]
    if T < 2:
        return elems
    npair = T // 2
    left = tuple(e[0:2 * npair:2] for e in elems)
    right = tuple(e[1:2 * npair:2] for e in elems)
    red = combine(left, right)                       # products of (2k, 2k+1)
    sc = associative_scan(combine, red)              # prefix at the odd positions 2k+1
    out = tuple(torch.empty_like(e) for e in elems)
    for o, e, s in zip(out, elems, sc):
        o[0] = e[0]
        o[1:2 * npair:2] = s
    if T > 2:
        nev = (T - 1) // 2                           # even positions 2, 4, ...: prefix(2k-1) x elem(2k)
        prev = tuple(s[:nev] for s in sc)
        cur = tuple(e[2:2 * nev + 1:2] for e in elems)
        ev = combine(prev, cur)
        for o, v in zip(out, ev):
            o[2:2 * nev + 1:2] = v
    return out


def _filter_combine(ei, ej):
    Ai, bi, Ci, etai, Ji = ei
    Aj, bj, Cj, etaj, Jj = ej
    n = Ai.shape[-1]
    eye = torch.eye(n, dtype=Ai.dtype, device=Ai.device)
    W = eye + Ci @ Jj
    rhs = torch.cat([Ai, (bi + (Ci @ etaj[..., None])[..., 0])[..., None], Ci], dim=-1)     # n x (2n+1)
    X = torch.linalg.solve(W, rhs)
    MAi, Mb, MCi = X[..., :n], X[..., n], X[..., n + 1:]
    rhs2 = torch.cat([(etaj - (Jj @ bi[..., None])[..., 0])[..., None], Jj @ Ai], dim=-1)   # n x (1+n)
    Y = torch.linalg.solve(W.transpose(-1, -2), rhs2)
    A = Aj @ MAi
    b = (Aj @ Mb[..., None])[..., 0] + bj
    C = Aj @ MCi @ Aj.transpose(-1, -2) + Cj
    eta = (Ai.transpose(-1, -2) @ Y[..., :1])[..., 0] + etai
    J = Ai.transpose(-1, -2) @ Y[..., 1:] + Ji
    return A, b, 0.5 * (C + C.transpose(-1, -2)), eta, 0.5 * (J + J.transpose(-1, -2))


def _affine_combine_rev(later, earlier):
    """Composition for the reverse scan: `later` = maps already composed from the end, `earlier` = the
    map of the step in front of them:  x_t = G_e (G_l x + c_l) + c_e."""
    Gl, cl = later
    Ge, ce = earlier
    return Ge @ Gl, (Ge @ cl[..., None])[..., 0] + ce


def kalman_sample_scan(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, w, w_last, return_filter=False):
    """Batched over sessions.  ys (N,T',p), mask (N,T'-1), zs (N,T'-1) int, m0 (n), S0 (n,n),
    A (K,n,n), B (K,n), Q (K,n,n), C (p,n), D (p), Rs (N,T',p), w (N,T'-1,n), w_last (N,n); float64.
    Returns xs (N,T',n) and the innovation-form marginal log-likelihood (N); with `return_filter`
    the filtered moments as kalman_filter emits them instead of xs (masked steps: the carry)."""
    N, Tp, p = ys.shape
    n = m0.shape[0]
    o = dict(dtype=torch.float64, device=ys.device)
    live = mask > 0                                               # (N, T'-1)
    # observation information of every frame
    CtR = C.T[None, None] / Rs[:, :, None, :]                     # (N,T',n,p)
    Jobs = CtR @ C                                                # (N,T',n,n)
    eobs = (CtR @ (ys - D)[..., None])[..., 0]                    # (N,T',n)
    eye = torch.eye(n, **o)
    zt = zs.long()
    lv = live[..., None, None].to(torch.float64)
    # elements, time first: [prior, step 0, ..., step T'-2]
    Ael = torch.where(live[..., None, None], A[zt], eye.expand(N, Tp - 1, n, n))
    bel = torch.where(live[..., None], B[zt], torch.zeros((), **o))
    Cel = Q[zt] * lv
    etael = eobs[:, :-1] * lv[..., 0]
    Jel = Jobs[:, :-1] * lv

    def tf(first, rest):   # (N, T'-1, ...) -> (T', N, ...) with the prior in front
        return torch.cat([first.expand(N, *first.shape)[None], rest.transpose(0, 1)], dim=0)
    elems = (tf(torch.zeros((n, n), **o), Ael), tf(m0, bel), tf(S0, Cel), tf(torch.zeros((n,), **o), etael),
             tf(torch.zeros((n, n), **o), Jel))
    _, m_pred, S_pred, _, _ = associative_scan(_filter_combine, elems)
    m_pred, S_pred = m_pred.transpose(0, 1), S_pred.transpose(0, 1)          # (N,T',n), (N,T',n,n): carry BEFORE y_t
    # conditioning on y_t (kalman.py:91-95) where the reference does it: unmasked steps and the last one
    upd = torch.cat([live, torch.ones((N, 1), dtype=torch.bool, device=ys.device)], dim=1)
    Sinv = torch.linalg.inv(S_pred)
    S_f = torch.linalg.inv(Sinv + Jobs)
    S_f = 0.5 * (S_f + S_f.transpose(-1, -2))
    m_f = (S_f @ ((Sinv @ m_pred[..., None])[..., 0] + eobs)[..., None])[..., 0]
    # innovation-form log-likelihood of the conditioned frames
    r = ys - D - (m_pred @ C.T)
    V = C @ S_pred @ C.T + torch.diag_embed(Rs)
    Lv = torch.linalg.cholesky(V)
    a = torch.linalg.solve_triangular(Lv, r[..., None], upper=False)[..., 0]
    lls = -0.5 * (a * a).sum(-1) - torch.log(torch.diagonal(Lv, dim1=-2, dim2=-1)).sum(-1) - 0.5 * p * math.log(2 * math.pi)
    ll = (lls * upd.to(torch.float64)).sum(1)
    if return_filter:
        return (torch.where(upd[..., None], m_f, m_pred), torch.where(upd[..., None, None], S_f, S_pred)), ll
    # backward sampler (kalman.py:123-156)
    x_last = m_f[:, -1] + (torch.linalg.cholesky(S_f[:, -1]) @ w_last[..., None])[..., 0]
    Qinv = torch.linalg.inv(Q)
    AtQ = (A.transpose(-1, -2) @ Qinv)[zt]                                    # (N,T'-1,n,n)
    Sfi = torch.linalg.inv(S_f[:, :-1])
    S_c = torch.linalg.inv(Sfi + AtQ @ A[zt])
    S_c = 0.5 * (S_c + S_c.transpose(-1, -2))
    Gm = S_c @ AtQ
    cm = (S_c @ ((Sfi @ m_f[:, :-1, :, None])[..., 0] - (AtQ @ B[zt][..., None])[..., 0])[..., None])[..., 0] \
        + (torch.linalg.cholesky(S_c) @ w[..., None])[..., 0]
    Gm = torch.where(live[..., None, None], Gm, eye.expand(N, Tp - 1, n, n))   # masked: x passes through
    cm = cm * lv[..., 0]
    Gs, cs = associative_scan(_affine_combine_rev, (Gm.transpose(0, 1), cm.transpose(0, 1)), reverse=True)
    xs_head = (Gs @ x_last[None, :, :, None])[..., 0] + cs                     # (T'-1, N, n)
    xs = torch.cat([xs_head.transpose(0, 1) * lv[..., 0], x_last[:, None]], dim=1)
    return xs, ll


def resample_latents_scan(key, *, Y, mask, v, h, z, s, Cd, sigmasq, Ab, Q, Gamma=None, draws=None,
                          return_info=False, **kwargs):
    """`gibbs.resample_latents` (gibbs.py:10-23) by the associative scan; same arguments and draws
    (`draws={'w': (N,T-L,n), 'w_last': (N,n)}` are required here: this variant generates none)."""
    if draws is None:
        raise ValueError("resample_latents_scan needs draws={'w', 'w_last'}")
    f64 = lambda t: torch.as_tensor(t, device=torch.as_tensor(Y).device).to(torch.float64)
    Y, mask, v, h, s, Cd, sigmasq, Ab, Q = (f64(t) for t in (Y, mask, v, h, s, Cd, sigmasq, Ab, Q))
    z = torch.as_tensor(z, device=Y.device)
    k, dk = Y.shape[-2:]
    D = Ab.shape[1]
    L = Ab.shape[2] // D
    Gamma = util.center_embedding(k, Y.device) if Gamma is None else f64(Gamma)
    Cdf = util.expand_Cd(Cd, Gamma, dk)                                        # gibbs.py:14
    ys = util.inverse_affine_transform(Y, v, h).reshape(*Y.shape[:-2], -1)     # :15
    A_, B_, Q_, C_ = util.ar_to_lds(Ab[..., :-1], Ab[..., -1], Q, Cdf[..., :-1])   # :16
    R = torch.repeat_interleave(s * sigmasq, dk, dim=-1)[:, L - 1:]            # :17
    n = D * L
    m0 = torch.zeros(n, dtype=torch.float64, device=Y.device)
    S0 = torch.eye(n, dtype=torch.float64, device=Y.device) * 10               # :18
    xs, ll = kalman_sample_scan(ys[:, L - 1:], mask[:, L - 1:-1], z, m0, S0, A_, B_, Q_, C_, Cdf[..., -1], R,
                                f64(draws["w"]), f64(draws["w_last"]))
    x = torch.cat([xs[:, 0, :-D].reshape(-1, L - 1, D), xs[:, :, -D:]], dim=1)    # :22
    return (x, {"loglik": ll}) if return_info else x


# ------------------------------------------------------------------ the reference's generic entry points
def _one(t, dev):
    return torch.as_tensor(t, device=dev).to(torch.float64)


def kalman_filter(ys, mask, zs, m0, S0, A, B, Q, C, D, Rs):
    """kalman.py:80-120 for ONE sequence of a generic switching LDS, by the associative scan:
    returns (filtered_ms (T',n), filtered_Ss (T',n,n)); masked steps emit the carry."""
    dev = torch.as_tensor(ys).device
    a = [_one(t, dev) for t in (ys, mask, m0, S0, A, B, Q, C, D, Rs)]
    ys, mask, m0, S0, A, B, Q, C, D, Rs = a
    n = m0.shape[0]
    z = torch.as_tensor(zs, device=dev)
    w = torch.zeros((1, ys.shape[0] - 1, n), dtype=torch.float64, device=dev)
    (ms, Ss), _ = kalman_sample_scan(ys[None], mask[None], z[None], m0, S0, A, B, Q, C, D, Rs[None], w,
                                     w[:, 0], return_filter=True)
    return ms[0], Ss[0]


def kalman_sample(rng, ys, mask, zs, m0, S0, A, B, Q, C, D, Rs, draws=None):
    """kalman.py:123-156 for ONE sequence: a sample of the states given the observations.  `rng`: int
    seed / (seed, offset) / torch.Generator as everywhere in this package; `draws={'w','w_last'}`
    injects the standard normals (T'-1, n) and (n)."""
    from . import distributions as dist
    dev = torch.as_tensor(ys).device
    ys, mask, m0, S0, A, B, Q, C, D, Rs = [_one(t, dev) for t in (ys, mask, m0, S0, A, B, Q, C, D, Rs)]
    n, Tp = m0.shape[0], ys.shape[0]
    if draws is not None:
        w, wl = _one(draws["w"], dev), _one(draws["w_last"], dev)
    elif dev.type == "cuda":
        seed = dist.seed_of(rng)
        w = dist.fill_normal((1, Tp - 1, n), seed, dist.STREAM_X, dev, 0)[0].double()
        wl = dist.fill_normal((1, n), seed, dist.STREAM_X_LAST, dev, 0)[0].double()
    else:
        raise ValueError("kalman_sample on CPU tensors needs injected draws")
    xs, _ = kalman_sample_scan(ys[None], mask[None], torch.as_tensor(zs, device=dev)[None], m0, S0, A, B, Q, C, D,
                               Rs[None], w[None], wl[None])
    return xs[0]
