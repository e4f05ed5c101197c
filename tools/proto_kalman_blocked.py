"""NumPy prototype for the next round's Kalman-filter lead: the bottom-up
factorisation  W = M^T M  (M lower triangular) and the inverse  L = M^-1  of kalman.cu, done in
8x8 blocks so that all O(n^3) work is 8x8x(4k) matrix products (DMMA m8n8k4 shaped) and only the
four diagonal blocks need scalar pivots.

    W_ij = sum_{k >= max(i,j)} M_ki^T M_kj        (block indices, M_ki = 0 for k < i)

  for k = nb-1 .. 0:
      M_kk  = bottom-up Cholesky of the (updated) diagonal block      scalar, 8 pivots
      Li_kk = M_kk^-1                                                 scalar, 8x8 triangular
      M_kj  = Li_kk^T W_kj              for j < k                     2 DMMA per block
      W_ij -= M_ki^T M_kj               for j <= i < k                2 DMMA per block
  L = M^-1 by blocks:  L_ii = Li_ii,  L_ij = -Li_ii sum_{j <= k < i} M_ik L_kj   (i > j)

Counts for n = 30 (nb = 4): 6 + 3 + 1 = 10 trailing blocks + 6 panel blocks + 16 inverse products
= 32 block products = 64 DMMA, against 870 DFMA-class warp instructions and 866 broadcast
shared-memory wavefronts in the lane-per-column version.  Run:  python tools/proto_kalman_blocked.py
"""
import numpy as np

B = 8


def chol_bottom_up(W):
    """M lower with W = M^T M (unblocked; the diagonal-block routine)."""
    n = W.shape[0]
    W = W.copy()
    M = np.zeros_like(W)
    for k in range(n - 1, -1, -1):
        M[k, k] = np.sqrt(W[k, k])
        M[k, :k] = W[k, :k] / M[k, k]
        W[:k, :k] -= np.outer(M[k, :k], M[k, :k])
    return M


def blocked(W):
    n = W.shape[0]
    nb = (n + B - 1) // B
    P = nb * B
    Wp = np.eye(P)
    Wp[:n, :n] = W                      # identity padding keeps the padded system SPD
    blk = lambda A, i, j: A[i * B:(i + 1) * B, j * B:(j + 1) * B]
    M = np.zeros((P, P))
    Li = [None] * nb
    products = 0
    for k in range(nb - 1, -1, -1):
        blk(M, k, k)[:] = chol_bottom_up(blk(Wp, k, k))
        Li[k] = np.linalg.inv(blk(M, k, k))          # 8x8 lower-triangular inverse (scalar code on the GPU)
        for j in range(k):
            blk(M, k, j)[:] = Li[k].T @ blk(Wp, k, j)
            products += 1
        for i in range(k):
            for j in range(i + 1):
                blk(Wp, i, j)[:] -= blk(M, k, i).T @ blk(M, k, j)
                if i != j:
                    blk(Wp, j, i)[:] = blk(Wp, i, j).T
                products += 1
    L = np.zeros((P, P))
    for i in range(nb):
        blk(L, i, i)[:] = Li[i]
        for j in range(i - 1, -1, -1):
            acc = np.zeros((B, B))
            for k in range(j, i):
                acc += blk(M, i, k) @ blk(L, k, j)
                products += 1
            blk(L, i, j)[:] = -Li[i] @ acc
            products += 1
    return M[:n, :n], L[:n, :n], products


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for n in (30, 21, 20, 8, 6):
        A = rng.standard_normal((n, n + 5))
        W = A @ A.T + n * np.eye(n)
        M, L, products = blocked(W)
        Mref = chol_bottom_up(W)
        e1 = np.abs(M - Mref).max()
        e2 = np.abs(M.T @ M - W).max() / np.abs(W).max()
        e3 = np.abs(L @ M - np.eye(n)).max()
        e4 = np.abs(L - np.linalg.cholesky(np.linalg.inv(W))).max()   # L = chol(S_cond), kalman.py:138
        print(f"n={n:2d} block products={products:3d}  |M-Mref|={e1:.1e} |M^T M-W|/|W|={e2:.1e} |L M-I|={e3:.1e} "
              f"|L-chol(W^-1)|={e4:.1e}")
