"""A NumPy stand-in for the small part of the JAX API the reference's hot path uses, so that the
reference's OWN, unmodified source files (keypoint_moseq/{gibbs,kalman,distributions,util}.py under
/root/reference) can be executed in the build container, where JAX and tensorflow-probability are
not installed.  Used only by tests/golden/make_reference_sweep.py to produce golden vectors.

What is the reference's and what is not:
  * every array expression, scan body, filter / sampler recursion, sufficient statistic and
    posterior formula that runs is the reference's code;
  * jit is the identity, vmap / lax.scan / lax.map / lax.cond are the obvious Python loops;
  * random numbers are INJECTED: jr.normal / jr.gamma return pre-drawn arrays, jr.categorical is the
    inverse CDF of one pre-drawn uniform, jr.multivariate_normal is mean + cholesky(cov) @ normal
    (JAX's default method), tfd.VonMises.sample is Best-Fisher rejection on pre-drawn uniforms --
    the same conventions the oracle states for these third-party samplers;
  * jr.dirichlet is the normalisation of pre-drawn gammas; the host-side np.random.binomial of
    transitions.py:29 is replaced (by the generator script) with a Bernoulli sum over pre-drawn
    uniforms; numba's private np.random.random stream inside sample_crp_tablecounts is the
    reference's own (seeded, and replayed as the injected uniforms);
  * tfd.MultivariateNormalFullCovariance.log_prob is restated from its definition.
"""
import sys
import types

import numpy as np
import scipy.linalg


class _AtIdx:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def set(self, v):
        out = np.array(self.arr, copy=True).view(SA)
        out[self.idx] = v
        return out


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self.arr, idx)


class SA(np.ndarray):
    """ndarray with the functional `.at[idx].set(v)` update."""
    @property
    def at(self):
        return _At(self)


def sa(x):
    return np.asarray(x).view(SA) if isinstance(x, np.ndarray) else x


class _NpProxy:
    def __init__(self, target):
        self._t = target

    def __getattr__(self, name):
        a = getattr(self._t, name)
        if isinstance(a, types.ModuleType):
            return _NpProxy(a)
        if callable(a) and not isinstance(a, type):
            def wrapped(*args, **kw):
                r = a(*args, **kw)
                if isinstance(r, tuple):
                    return tuple(sa(x) for x in r)
                return sa(r)
            return wrapped
        return a


class Key:
    """Opaque stand-in for a PRNG key; draws are looked up in CTX, not derived from it."""
    def __iter__(self):
        raise TypeError("Key is not iterable")


class KeyList(list):
    pass


class Ctx:
    draws = {}       # name -> array (leading axes indexed by the active vmap / scan indices)
    vmap = []        # stack of active vmap indices
    scan_t = None    # index of the active lax.scan iteration


def _take(name):
    a = Ctx.draws[name]
    for i in Ctx.vmap:
        a = a[i]
    return a


# ------------------------------------------------------------------ jax.random
def split(key, num=2):
    return KeyList(Key() for _ in range(num))


def normal(key, shape=(), dtype=None):
    shape = tuple(shape)
    for name in ("normal", "normal2"):
        if name in Ctx.draws:
            a = _take(name)
            if tuple(a.shape) == shape:
                return sa(np.array(a, dtype=np.float64))
    raise KeyError(f"no injected normal draw of shape {shape}")


def multivariate_normal(key, mean, cov):
    w = _take("mvn")
    return sa(np.asarray(mean) + np.linalg.cholesky(np.asarray(cov)) @ w)


def gamma(key, a, dtype=None):
    g = _take("gamma")
    assert np.broadcast(g, np.asarray(a)).shape == g.shape
    return sa(np.array(g, dtype=np.float64))


def dirichlet(key, alpha):
    """jr.dirichlet = normalised Gamma(alpha, 1) variates; the injected gammas are consumed in call order."""
    g = np.asarray(Ctx.draws["dirichlet_g"].pop(0), dtype=np.float64)
    assert g.shape == np.asarray(alpha).shape, (g.shape, np.asarray(alpha).shape)
    return sa(g / g.sum(-1, keepdims=True))


def bernoulli_sum_binomial(u_thin, rho_all):
    """Stand-in for the host-side np.random.binomial(n, p) of transitions.py:29 (called once, on the
    non-empty diagonal cells): the entry for state i is the Bernoulli sum over the injected uniforms
    u_thin[i, :n] — the convention the oracle and the CUDA kernel state for this draw.  The state
    of an entry is recovered from its success probability (rho_all[i], distinct per state)."""
    def binomial(n, p, size=None):
        n, p = np.atleast_1d(np.asarray(n)), np.atleast_1d(np.asarray(p))
        rows = [int(np.argmin(np.abs(rho_all - pp))) for pp in p]
        return np.array([(u_thin[r, :int(nn)] < pp).sum() for r, nn, pp in zip(rows, n, p)], dtype=np.int64)
    return binomial


def categorical(key, logits):
    u = _take("uniform")[Ctx.scan_t]
    w = np.exp(np.asarray(logits, dtype=np.float64))
    c = np.cumsum(w)
    return int(min(np.searchsorted(c, u * c[-1], side="right"), len(w) - 1))


# ------------------------------------------------------------------ jax.lax / transforms
def _leaves_len(xs):
    if isinstance(xs, (tuple, list)) and not isinstance(xs, KeyList):
        return _leaves_len(xs[0])
    return len(xs)


def _index(xs, i):
    if isinstance(xs, tuple):
        return tuple(_index(x, i) for x in xs)
    return xs[i]


def _stack(items):
    first = items[0]
    if isinstance(first, tuple):
        return tuple(_stack([it[j] for it in items]) for j in range(len(first)))
    return sa(np.stack([np.asarray(it) for it in items]))


def scan(f, init, xs, length=None, reverse=False):
    n = _leaves_len(xs)
    order = range(n - 1, -1, -1) if reverse else range(n)
    carry, ys = init, [None] * n
    saved = Ctx.scan_t
    for t in order:
        Ctx.scan_t = t
        carry, ys[t] = f(carry, _index(xs, t))
    Ctx.scan_t = saved
    return carry, _stack(ys)


def cond(pred, true_fun, false_fun, *operands):
    return true_fun(*operands) if bool(pred) else false_fun(*operands)


def lax_map(f, xs):
    return _stack([f(_index(xs, i)) for i in range(_leaves_len(xs))])


def vmap(f, in_axes=0, out_axes=0):
    def g(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = next(len(a) for a, ax in zip(args, axes) if ax == 0)
        outs = []
        for i in range(n):
            Ctx.vmap.append(i)
            try:
                outs.append(f(*[(a[i] if ax == 0 else a) for a, ax in zip(args, axes)]))
            finally:
                Ctx.vmap.pop()
        return _stack(outs)
    return g


def jit(f=None, **kw):
    return f if f is not None else (lambda g: g)


# ------------------------------------------------------------------ tfd
class MultivariateNormalFullCovariance:
    def __init__(self, loc, covariance_matrix):
        self.loc, self.cov = np.asarray(loc), np.asarray(covariance_matrix)

    def log_prob(self, x):
        d = np.asarray(x) - self.loc
        L = np.linalg.cholesky(self.cov)
        y = scipy.linalg.solve_triangular(L, d.reshape(-1, d.shape[-1]).T, lower=True)
        q = (y * y).sum(0).reshape(d.shape[:-1])
        n = d.shape[-1]
        return sa(-0.5 * q - np.log(np.diag(L)).sum() - 0.5 * n * np.log(2 * np.pi))


class VonMises:
    sampler = None   # set by the generator script: (theta, kappa, u) -> sample

    def __init__(self, loc, concentration):
        self.loc, self.k = np.asarray(loc), np.asarray(concentration)

    def sample(self, seed=None):
        return sa(VonMises.sampler(self.loc, self.k, _take("vonmises")))


def import_reference(*modules, ref="/root/reference"):
    """Import modules of the REFERENCE's `keypoint_moseq` (a namespace package: it has no
    __init__.py) although this repository ships a regular package of the same name (the drop-in
    shim), which would otherwise win: the repository root and the working directory are taken off
    sys.path while the reference is imported, and cached modules of that name are dropped."""
    import importlib
    import os
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    saved = list(sys.path)
    sys.path[:] = [ref] + [p for p in saved if p not in ("", ".", ref) and os.path.abspath(p) != here
                           and os.path.abspath(p) != os.getcwd()]
    for name in [m for m in sys.modules if m == "keypoint_moseq" or m.startswith("keypoint_moseq.")]:
        del sys.modules[name]
    try:
        out = [importlib.import_module("keypoint_moseq." + m) for m in modules]
        for mod in out:
            assert mod.__file__.startswith(ref), mod.__file__
    finally:
        sys.path[:] = saved
    return out if len(out) > 1 else out[0]


def install():
    """Register the stand-ins as `jax`, `jax.numpy`, ... in sys.modules."""
    jnp = _NpProxy(np)
    lax = types.SimpleNamespace(scan=scan, cond=cond, map=lax_map)
    jr = types.SimpleNamespace(split=split, normal=normal, multivariate_normal=multivariate_normal,
                               gamma=gamma, categorical=categorical, dirichlet=dirichlet,
                               PRNGKey=lambda seed: Key())
    jscipy = types.SimpleNamespace(linalg=types.SimpleNamespace(
        solve_triangular=lambda a, b, lower=False, **k: sa(scipy.linalg.solve_triangular(a, b, lower=lower))))
    config = types.SimpleNamespace(update=lambda *a, **k: None)
    jax = types.ModuleType("jax")
    jax.__dict__.update(jit=jit, vmap=vmap, lax=lax, random=jr, numpy=jnp, scipy=jscipy, config=config)
    tfd = types.SimpleNamespace(MultivariateNormalFullCovariance=MultivariateNormalFullCovariance, VonMises=VonMises)
    sys.modules["jax"] = jax
    sys.modules["jax.numpy"] = jnp
    sys.modules["jax.random"] = jr
    sys.modules["jax.config"] = types.SimpleNamespace(config=config)
    sys.modules["jax.scipy"] = jscipy
    tfp_jax = types.SimpleNamespace(distributions=tfd)
    tfp_sub = types.SimpleNamespace(jax=tfp_jax)
    sys.modules["tensorflow_probability"] = types.SimpleNamespace(substrates=tfp_sub)
    sys.modules["tensorflow_probability.substrates"] = tfp_sub
    sys.modules["tensorflow_probability.substrates.jax"] = tfp_jax
    sys.modules["tensorflow_probability.substrates.jax.distributions"] = tfd
    return jax
