"""kpms_ar_loglik_static: the emission log-likelihood for callers that may not allocate.

Some users of this library never get to call cudaMalloc: an XLA custom-call handler that
wraps these kernels for the JAX reference is given its operand and result buffers by XLA and
must leave the device's memory alone (XLA's own pool normally owns it already).  For them the
library keeps the small weight scratch of this kernel in a fixed array of its own.  This test
drives that entry point through the C-ABI exactly as such a caller would — inputs and the output
buffer only — and checks it against the regular entry point (which takes a caller workspace),
twice in a row with different problem sizes so that whatever the first call left in the shared
array cannot leak into the second.

The file name sorts last on purpose: if this test fails, `pytest -x` has already run the rest.
"""
import ctypes

import numpy as np
import pytest

import kpms_helpers as H

TOL = 2e-5   # same relative tolerance as tests/test_gpu_parity.py::test_ar_loglik_parity


def _both(lib, x, Ab, Q, impl):
    import torch
    from keypointmoseq_b200 import _lib
    N, T, D = x.shape
    K, L = Ab.shape[0], Ab.shape[2] // Ab.shape[1]
    st = _lib.stream_ptr(x.device)
    want = torch.empty((N, T - L, K), dtype=torch.float32, device=x.device)
    nbytes = lib.kpms_ar_loglik_workspace_bytes(K, D, L)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    _lib.check(lib.kpms_ar_loglik(_lib.ptr(x), _lib.ptr(Ab), _lib.ptr(Q), N, T, D, L, K, _lib.ptr(want),
                                  _lib.ptr(ws), nbytes, impl, st), "kpms_ar_loglik")
    got = torch.full((N, T - L, K), float("nan"), dtype=torch.float32, device=x.device)
    _lib.check(lib.kpms_ar_loglik_static(_lib.ptr(x), _lib.ptr(Ab), _lib.ptr(Q), N, T, D, L, K, _lib.ptr(got),
                                         impl, st), "kpms_ar_loglik_static")
    torch.cuda.synchronize()
    return got.cpu().numpy().astype(np.float64), want.cpu().numpy().astype(np.float64)


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_workspace_entry_point_twice():
    import torch
    from keypointmoseq_b200 import _lib
    lib = _lib.load()
    # two back-to-back calls: a large state count first (fills most of the shared array),
    # then a small, different problem
    for seed, K, T in [(5, 100, 90), (6, 7, 130)]:
        d, s, p, _ = H.problem(seed=seed, N=3, T=T, D=10, L=3, K=K, ragged=False)
        x = torch.as_tensor(s['x'], dtype=torch.float32, device="cuda").contiguous()
        Ab = torch.as_tensor(p['Ab'], dtype=torch.float64, device="cuda").contiguous()
        Q = torch.as_tensor(p['Q'], dtype=torch.float64, device="cuda").contiguous()
        for impl in (1, 0):
            got, want = _both(lib, x, Ab, Q, impl)
            assert np.all(np.isfinite(got))
            err = np.abs(got - want) / np.maximum(1.0, np.abs(want))
            assert err.max() < TOL, (K, impl, err.max())


def test_static_scratch_entry_point_refuses_oversize_problems():
    """Beyond the built-in capacity the call must say so and launch nothing (no GPU needed:
    the check comes before any CUDA call)."""
    from keypointmoseq_b200 import build, _lib
    build.build()
    lib = _lib.load()
    one = ctypes.c_void_p(256)
    rc = lib.kpms_ar_loglik_static(one, one, one, 1, 50, 10, 3, 200, one, 1, None)   # 200*10*44 floats > 65536
    assert rc == -6
    assert b"too large" in lib.kpms_status_string(rc)
    assert lib.kpms_ar_loglik_static(None, one, one, 1, 50, 10, 3, 4, one, 1, None) == -1
