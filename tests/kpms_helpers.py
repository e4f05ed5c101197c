"""Shared fixtures for the parity tests: seeded problems, injected draws, oracle calls."""
import numpy as np

from keypointmoseq_b200.synthetic import default_hypparams, simulate
from oracle import kpms_oracle as O


def problem(seed=0, N=3, T=80, k=8, dk=2, D=10, L=3, K=6, ragged=True):
    """Float32-rounded data/states (as float64 for the oracle) + float64 params."""
    data, st, pr, Gamma = simulate(seed, N, T, k=k, dk=dk, D=D, L=L, K=K, ragged=ragged, dtype=np.float32)
    d64 = {a: b.astype(np.float64) for a, b in data.items()}
    s64 = {a: (b.astype(np.float64) if a != 'z' else b) for a, b in st.items()}
    return d64, s64, pr, Gamma


def f32r(a):
    """Round to float32 and return as float64 (what the kernels see)."""
    return np.asarray(a, dtype=np.float32).astype(np.float64)


def sweep_draws(rng, N, T, k, dk, D, L, K, rounds=8):
    n = D * L
    return dict(
        z_u=f32r(rng.random((N, T - L))),
        x_w=f32r(rng.standard_normal((N, T - L, n))),
        x_w_last=f32r(rng.standard_normal((N, n))),
        h_u=f32r(rng.random((N, T, rounds, 3)) * 0.999 + 0.0005),
        v_w=f32r(rng.standard_normal((N, T - 1, dk))),
        v_w_last=f32r(rng.standard_normal((N, dk))),
    )


def hyp(D, L, K):
    trans, ar, obs, tr = default_hypparams(D, L, K)
    out = {}
    for d in (trans, ar, obs, tr):
        out.update(d)
    return out


def to_torch(d, device="cuda"):
    import torch
    out = {}
    for key, val in d.items():
        if isinstance(val, np.ndarray):
            if val.dtype.kind == 'i':
                out[key] = torch.as_tensor(val.astype(np.int32), device=device)
            else:
                out[key] = torch.as_tensor(val, device=device)
        else:
            out[key] = val
    return out
