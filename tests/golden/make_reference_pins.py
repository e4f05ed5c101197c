"""Golden vectors produced by RUNNING the reference's own code (run in the build container only).

JAX / tensorflow-probability are not installed, so most of the reference cannot be executed.  Two
functions on the hot path are plain numba, though, and run as they are once the module-level
`import jax ...` lines are satisfied by inert stand-ins (nothing of the stand-ins is ever called
by these two functions):

  keypoint_moseq/transitions.py:7-21   sample_crp_tablecounts   (row a11)
  keypoint_moseq/util.py:178-189       count_transitions        (row a10)

sample_crp_tablecounts draws `np.random.random(tot)` from numba's private Mersenne Twister.  Seeded
inside numba with s, that stream is the same MT19937 double stream as NumPy's legacy
`RandomState(s).random_sample(tot)`, which is what the oracle and the CUDA kernel are then given as
injected uniforms; the fixture stores inputs, the seed and the reference's outputs.

    python tests/golden/make_reference_pins.py      # writes tests/golden/reference_pins.npz
"""
import importlib
import os
import sys
import types

import numpy as np

sys.dont_write_bytecode = True
REF = "/root/reference"


class _Inert:
    """Import-time stand-in: any attribute, callable, usable as a decorator."""
    def __getattr__(self, k):
        return _Inert()

    def __call__(self, *a, **k):
        if len(a) == 1 and callable(a[0]) and not k:
            return a[0]
        return _Inert()


def _load_reference():
    inert = _Inert()
    jax = types.ModuleType("jax")
    jax.__dict__.update(jit=lambda f=None, **k: (f if f is not None else (lambda g: g)), vmap=inert, lax=inert,
                        random=inert, numpy=inert, config=inert)
    sys.modules["jax"] = jax
    for name in ["jax.config", "jax.numpy", "jax.random", "jax.scipy", "jax.scipy.special",
                 "tensorflow_probability", "tensorflow_probability.substrates",
                 "tensorflow_probability.substrates.jax",
                 "tensorflow_probability.substrates.jax.distributions"]:
        sys.modules[name] = inert
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from jax_numpy_shim import import_reference   # the repository's own `keypoint_moseq` shim must not win
    T, U = import_reference("transitions", "util", ref=REF)
    return T.sample_crp_tablecounts, U.count_transitions


def main():
    from numba import njit
    ref_crp, ref_count = _load_reference()

    @njit
    def seed_numba(s):
        np.random.seed(s)

    out = {}
    rng = np.random.default_rng(20240611)
    crp_cases = [(5, 0.7, 40, 11), (12, 5.7, 300, 12), (30, 100.0, 2000, 13), (100, 5.7, 400, 14)]
    for ci, (K, alpha, cmax, seed) in enumerate(crp_cases):
        customers = rng.integers(0, cmax, size=(K, K)).astype(np.int64)
        customers[rng.random((K, K)) < 0.3] = 0                      # empty cells
        customers[np.arange(K), np.arange(K)] += 5 * cmax            # sticky diagonal
        betas = rng.dirichlet(np.ones(K))
        seed_numba(seed)
        m_ref = ref_crp(float(alpha), customers, betas)
        out[f"crp{ci}_seed"] = np.int64(seed)     # uniforms = np.random.RandomState(seed).random_sample(tot)
        out[f"crp{ci}_alpha"] = np.float64(alpha)
        out[f"crp{ci}_customers"] = customers
        out[f"crp{ci}_betas"] = betas
        out[f"crp{ci}_m"] = np.asarray(m_ref, dtype=np.int64)
    out["n_crp"] = np.int64(len(crp_cases))

    cnt_cases = [(3, 40, 4, False), (6, 700, 25, True), (4, 3000, 100, True)]
    for ci, (N, T, K, ragged) in enumerate(cnt_cases):
        z = rng.integers(0, K, size=(N, T)).astype(np.int64)
        mask = np.ones((N, T))
        if ragged:
            for n in range(N):
                mask[n, rng.integers(T // 2, T):] = 0
            mask[0, 5:9] = 0                                          # an interior hole
        c_ref = ref_count(K, mask, z)
        out[f"cnt{ci}_K"] = np.int64(K)
        out[f"cnt{ci}_z"] = z
        out[f"cnt{ci}_mask"] = mask
        out[f"cnt{ci}_counts"] = np.asarray(c_ref, dtype=np.float64)
    out["n_cnt"] = np.int64(len(cnt_cases))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_pins.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: getattr(v, "shape", ()) for k, v in out.items() if k.endswith("_m") or k.endswith("_counts")})


if __name__ == "__main__":
    main()
