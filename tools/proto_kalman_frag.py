"""Lane-level NumPy emulation of one step of the blocked (8x8, DMMA m8n8k4) Kalman filter kernel
(kalman_forward_blk_kernel in keypointmoseq_b200/csrc/kalman.cu): every register is an array over
the 32 lanes, shared memory is a flat array, the FP64 tensor-core instruction is emulated from its
fragment layout.  It pins the index arithmetic of the kernel against a dense NumPy step.

    python tools/proto_kalman_frag.py

Fragment layouts of mma.sync.m8n8k4.f64 (lane = 4 g + qd):
    A (8x4, row):  lane holds A[g][qd]         B (4x8, col):  lane holds B[qd][g]
    C (8x8):       lane holds C[g][2 qd], C[g][2 qd + 1]
Buffer fragments of a row-major matrix X with row stride SA (SA = 4 mod 8 keeps all three conflict-free):
    K-fragment  X[r0 + qd][c0 + g]   (A operand for X^T, B operand for X)
    R-fragment  X[r0 + g][c0 + qd]   (A operand for X)
    C-fragment  X[r0 + g][c0 + 2 qd + e]
"""
import numpy as np

LANES = np.arange(32)
G_, QD = LANES >> 2, LANES & 3


def dmma(acc, a, b):
    """acc (32,2) += A @ B with A[g][k] = a[4g+k], B[k][n] = b[4n+k]."""
    A = a.reshape(8, 4)
    B = b.reshape(8, 4).T
    C = A @ B
    acc[:, 0] += C[G_, 2 * QD]
    acc[:, 1] += C[G_, 2 * QD + 1]


class Buf:
    def __init__(self, rows, stride):
        self.m = np.zeros((rows, stride))

    def kfrag(self, r0, c0):
        return self.m[r0 + QD, c0 + G_].copy()

    def rfrag(self, r0, c0):
        return self.m[r0 + G_, c0 + QD].copy()

    def cstore(self, r0, c0, acc):
        self.m[r0 + G_, c0 + 2 * QD] = acc[:, 0]
        self.m[r0 + G_, c0 + 2 * QD + 1] = acc[:, 1]

    def cload(self, r0, c0):
        return np.stack([self.m[r0 + G_, c0 + 2 * QD], self.m[r0 + G_, c0 + 2 * QD + 1]], axis=1)


def blk(i, j):
    return i * (i + 1) // 2 + j


def step(NN, DD, Lam_p, eta_p, Gz, J, e_obs, FD, f, Qi, qb):
    """One live filter step.  Returns L (NN,NN), eta_c, Lam_p', eta_p', prod(1/M_dd), u^T u."""
    OFF = NN - DD
    NBK = (NN + 1 + 7) // 8
    P, SA = 8 * NBK, 8 * NBK + 4
    CB0 = OFF // 8
    # ---- carry and state table in C-fragment registers; row NN of the augmented matrix carries eta
    Caug = np.zeros((P, P))
    Caug[:NN, :NN] = Lam_p
    Caug[NN, :NN] = eta_p
    Gaug = np.zeros((P, P))
    Gaug[:NN, :NN] = Gz
    for d in range(NN, P):
        Gaug[d, d] = 1.0                                    # identity padding keeps W SPD
    frag = lambda M, i, j: np.stack([M[8 * i + G_, 8 * j + 2 * QD], M[8 * i + G_, 8 * j + 2 * QD + 1]], axis=1)
    C = {(i, j): frag(Caug, i, j) for i in range(NBK) for j in range(i + 1)}
    Gf = {(i, j): frag(Gaug, i, j) for i in range(NBK) for j in range(i + 1)}
    # ---- S1: W = Lam_p + G + J on rows < NN, padding rows from the table only; eta_c
    W = {}
    for (i, j) in C:
        w = np.zeros((32, 2))
        for e in range(2):
            row, col = 8 * i + G_, 8 * j + 2 * QD + e
            inside = (row < NN) & (col < NN)
            w[:, e] = np.where(inside, C[(i, j)][:, e], 0.0) + Gf[(i, j)][:, e]
            jm = inside & (row >= OFF) & (col >= OFF) & (col <= row)
            a, b = np.clip(row - OFF, 0, DD - 1), np.clip(col - OFF, 0, DD - 1)
            Jv = np.where(jm, J[a, b], 0.0)
            w[:, e] += Jv
        W[(i, j)] = w
    hv = np.zeros(P)
    eta_c = np.zeros(NN)
    ib = NN // 8                                            # block row of the eta row
    for j in range(ib + 1):
        for e in range(2):
            col = 8 * j + 2 * QD + e
            sel = (8 * ib + G_ == NN) & (col < NN)
            ec = C[(ib, j)][:, e] + np.where(col >= OFF, e_obs[np.clip(col - OFF, 0, DD - 1)], 0.0)
            eta_c[col[sel]] = ec[sel]
            hv[col[sel]] = ec[sel] - f[col[sel]]
    # ---- phase A: bottom-up blocked Cholesky W = M^T M, left-looking; M off-diagonal blocks and the
    #      inverses of the diagonal blocks go to the buffer
    buf = Buf(P, SA)
    pd = 1.0
    for k in range(NBK - 1, -1, -1):
        acc = {j: W[(k, j)].copy() for j in range(k + 1)}
        for kp in range(k + 1, NBK):
            for half in range(2):
                a = -buf.kfrag(8 * kp + 4 * half, 8 * k)              # -(M_k'k)^T chunk
                for j in range(k + 1):
                    dmma(acc[j], a, buf.kfrag(8 * kp + 4 * half, 8 * j))
        # block row k by scalar pivots, lane c owns column c of the strip (rows 8k..8k+7, columns <= 8k+7):
        # diagonal block and panel are factored together, no inverse of the diagonal block is needed here
        for j in range(k + 1):
            buf.cstore(8 * k, 8 * j, acc[j])
        ncol = 8 * (k + 1)
        R = np.zeros((32, 8))                                          # R[c][i] = strip element (8k+i, c)
        for c in range(ncol):
            R[c, :] = buf.m[8 * k:8 * k + 8, c]
        Mrow = np.zeros((32, 8))
        piv = R[8 * k + 7, 7]
        for d in range(7, -1, -1):
            r = 1.0 / piv
            rs = np.sqrt(r)
            pd *= rs
            t = R[:, d] * r                                            # W[d][c] / piv, every lane its own column
            if d > 0:
                cl = 8 * k + d - 1                                     # lane that owns the next pivot
                piv = R[cl, d - 1] - t[cl] * R[cl, d]
            Mrow[:, d] = R[:, d] * rs
            for i in range(d):                                         # rows above the pivot, all columns
                R[:, i] -= t[8 * k + i] * R[:, d]
        for c in range(ncol):
            for d in range(8):
                buf.m[8 * k + d, c] = Mrow[c, d] if c <= 8 * k + d else 0.0
    # ---- inverses of the four diagonal blocks in one pass: lane = 8 b + c computes column c of M_bb^-1
    Lis = []
    for b in range(NBK):
        Mb = buf.m[8 * b:8 * b + 8, 8 * b:8 * b + 8].copy()
        Li = np.zeros((8, 8))
        for c in range(8):
            for i in range(8):
                acc_ = 1.0 if i == c else 0.0
                for q in range(c, i):
                    acc_ -= Mb[i, q] * Li[q, c]
                Li[i, c] = acc_ / Mb[i, i] if i >= c else 0.0
        Lis.append(Li)
    for b in range(NBK):
        buf.m[8 * b:8 * b + 8, 8 * b:8 * b + 8] = Lis[b]
    # ---- phase B: L_ij = -(sum_{k=j+1..i} L_ik M_kj) Li_jj.  First every off-diagonal block of M is
    #      scaled on the right, N_kj = M_kj Li_jj (six independent products, in place); then
    #      L_ij = -sum_k L_ik N_kj needs one product per term and no layout change inside the chain.
    #      Blocks at the same distance from the diagonal are independent.  N stays in the lower blocks
    #      until the end: an L block is parked, untransposed, in the free upper block (j, i).
    Nn = {}
    for k in range(1, NBK):
        for j in range(k):
            n_ = np.zeros((32, 2))
            for half in range(2):
                dmma(n_, buf.rfrag(8 * k, 8 * j + 4 * half), buf.kfrag(8 * j + 4 * half, 8 * j))
            Nn[(k, j)] = n_
    for (k, j), n_ in Nn.items():
        buf.cstore(8 * k, 8 * j, n_)
    Lreg = {}
    for sd in range(1, NBK):
        for i in range(sd, NBK):
            j = i - sd
            l = np.zeros((32, 2))
            for k in range(j + 1, i + 1):
                for half in range(2):
                    a = buf.rfrag(8 * i, 8 * i + 4 * half) if k == i else buf.rfrag(8 * k, 8 * i + 4 * half)
                    dmma(l, -a, buf.kfrag(8 * k + 4 * half, 8 * j))
            Lreg[(i, j)] = l
        for i in range(sd, NBK):
            buf.cstore(8 * (i - sd), 8 * i, Lreg[(i, i - sd)])
    for (i, j), l in Lreg.items():
        buf.cstore(8 * i, 8 * j, l)
    L = np.tril(buf.m[:NN, :NN])                                    # the upper blocks hold parked copies
    # ---- phase C: T = L^T [F | h] (dense column blocks CB0..NBK-1), Z = T^T T
    Tb = Buf(P, 8 * (NBK - CB0) + 4)
    KQ = 2 * NBK
    Tacc = {(rb, cb): np.zeros((32, 2)) for rb in range(NBK) for cb in range(NBK - CB0)}
    for kq in range(KQ):
        q = 4 * kq + QD
        bF = {}
        for cb in range(NBK - CB0):
            cf = 8 * (cb + CB0) + G_
            qq = np.minimum(q, NN - 1)
            fd = FD[qq, np.clip(cf - OFF, 0, DD - 1)]
            v = np.where(cf < OFF, np.where(q == DD + cf, 100.0, 0.0),
                         np.where(cf < NN, fd, np.where(cf == NN, hv[np.minimum(q, P - 1)], 0.0)))
            bF[cb] = np.where(q < NN, v, 0.0)
        for rb in range(NBK):
            if kq < 2 * rb:
                continue
            a = buf.kfrag(4 * kq, 8 * rb)                              # (L^T)[8rb+g][4kq+qd] = L[4kq+qd][8rb+g]
            a = np.where((q < NN) & (8 * rb + G_ <= q), a, 0.0)
            for cb in range(NBK - CB0):
                dmma(Tacc[(rb, cb)], a, bF[cb])
    for (rb, cb), t in Tacc.items():
        Tb.cstore(8 * rb, 8 * cb, t)
    Tk = {}
    for kq in range(KQ):
        r = 4 * kq + QD
        for b in range(NBK):
            cf = 8 * b + G_
            if b < CB0:
                src = buf.m[np.minimum(DD + cf, P - 1), np.minimum(r, P - 1)]
                Tk[(kq, b)] = np.where((r <= DD + cf) & (DD + cf < NN), 100.0 * src, 0.0)
            else:
                Tk[(kq, b)] = Tb.m[r, cf - 8 * CB0].copy()
    Z = {(ib_, jb): np.zeros((32, 2)) for ib_ in range(NBK) for jb in range(ib_ + 1)}
    for kq in range(KQ):
        for (ib_, jb) in Z:
            dmma(Z[(ib_, jb)], Tk[(kq, ib_)], Tk[(kq, jb)])
    # ---- new carry: Lam_p' = Qinv_ - Z, eta_p' = qb_ + Z[NN][:]
    Qaug = np.zeros((P, P))
    Qaug[:OFF, :OFF] = 100.0 * np.eye(OFF)
    Qaug[OFF:NN, OFF:NN] = Qi
    Qaug[NN, OFF:NN] = qb
    Cn = np.zeros((P, P))
    for (i, j), z in Z.items():
        for e in range(2):
            row, col = 8 * i + G_, 8 * j + 2 * QD + e
            val = np.where(row == NN, Qaug[row, col] + z[:, e], Qaug[row, col] - z[:, e])
            Cn[row, col] = val
    Lam_new = np.tril(Cn[:NN, :NN])
    Lam_new = Lam_new + np.tril(Lam_new, -1).T
    uu = None
    for (i, j), z in Z.items():
        for e in range(2):
            row, col = 8 * i + G_, 8 * j + 2 * QD + e
            hit = (row == NN) & (col == NN)
            if hit.any():
                uu = float(z[hit, e][0])
    return L, eta_c, Lam_new, Cn[NN, :NN].copy(), pd, uu


def dense_step(NN, DD, Lam_p, eta_p, Gz, J, e_obs, FD, f, Qi, qb):
    OFF = NN - DD
    W = Lam_p + Gz
    W[OFF:, OFF:] += J
    eta_c = eta_p.copy()
    eta_c[OFF:] += e_obs
    S = np.linalg.inv(W)
    L = np.linalg.cholesky(S)
    B = np.zeros((NN, NN + 1))
    for c in range(OFF):
        B[DD + c, c] = 100.0
    B[:, OFF:NN] = FD
    B[:, NN] = eta_c - f
    T = L.T @ B
    Z = T.T @ T
    Qinv = np.zeros((NN, NN))
    Qinv[:OFF, :OFF] = 100.0 * np.eye(OFF)
    Qinv[OFF:, OFF:] = Qi
    Lam_new = Qinv - Z[:NN, :NN]
    eta_new = np.concatenate([np.zeros(OFF), qb]) + Z[NN, :NN]
    M = np.linalg.inv(L)
    return L, eta_c, Lam_new, eta_new, np.prod(1.0 / np.diag(M)), Z[NN, NN]


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for NN, DD in [(30, 10), (21, 7), (20, 10), (10, 10), (15, 5), (8, 4), (6, 3), (4, 2)]:
        OFF = NN - DD
        A = rng.standard_normal((NN, NN + 3))
        Lam_p = A @ A.T / NN + np.eye(NN)
        eta_p = rng.standard_normal(NN)
        Fm = rng.standard_normal((DD, NN)) * 0.5                  # D x n AR block
        Qi = np.linalg.inv(0.05 * np.eye(DD) + 0.01 * np.cov(rng.standard_normal((DD, 40))))
        Gz = Fm.T @ Qi @ Fm
        Gz[DD:, DD:] += 100.0 * np.eye(OFF) if OFF > 0 else 0.0
        FD = Fm.T @ Qi                                             # n x D: last D columns of A^T Qinv_
        b = rng.standard_normal(DD) * 0.1
        f, qb = FD @ b, Qi @ b
        Cm = rng.standard_normal((12, DD))
        J = Cm.T @ Cm
        e_obs = rng.standard_normal(DD)
        got = step(NN, DD, Lam_p, eta_p, Gz, J, e_obs, FD, f, Qi, qb)
        want = dense_step(NN, DD, Lam_p.copy(), eta_p, Gz, J, e_obs, FD, f, Qi, qb)
        names = ["L", "eta_c", "Lam_p'", "eta_p'", "prod 1/M_dd", "u^T u"]
        errs = [float(np.abs(np.asarray(g) - np.asarray(w)).max() / (np.abs(np.asarray(w)).max() + 1e-300)) for g, w in zip(got, want)]
        print(f"NN={NN:2d} DD={DD:2d} " + "  ".join(f"{n} {e:.1e}" for n, e in zip(names, errs)))
        assert max(errs) < 1e-10, (NN, DD, errs)
    print("ok")
