"""ctypes binding of libkpms_b200.so (the C-ABI declared in include/kpms_b200.h).

There is no fallback: if the shared library is missing or a call fails, an exception
is raised.  Tensors cross the boundary as raw device pointers; torch only provides
device memory and the current CUDA stream.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkpms_b200.so")

c_int, c_size_t, c_void_p, c_double = ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_double
c_int64, c_uint64 = ctypes.c_int64, ctypes.c_uint64
P = c_void_p

# name -> (restype, argtypes); mirrors include/kpms_b200.h one to one
SIGNATURES = {
    "kpms_version": (c_int, []),
    "kpms_status_string": (ctypes.c_char_p, [c_int]),
    "kpms_launch_count": (c_uint64, []),
    "kpms_set_option": (c_int, [ctypes.c_char_p, c_int]),
    "kpms_get_option": (c_int, [ctypes.c_char_p, ctypes.POINTER(c_int)]),
    "kpms_fill_uniform_at": (c_int, [P, c_size_t, c_uint64, c_uint64, c_uint64, P]),
    "kpms_fill_normal_at": (c_int, [P, c_size_t, c_uint64, c_uint64, c_uint64, P]),
    "kpms_fill_gamma_at": (c_int, [P, c_size_t, c_uint64, c_double, c_uint64, c_uint64, P]),
    "kpms_fill_uniform": (c_int, [P, c_size_t, c_uint64, c_uint64, P]),
    "kpms_fill_normal": (c_int, [P, c_size_t, c_uint64, c_uint64, P]),
    "kpms_fill_gamma": (c_int, [P, c_size_t, c_double, c_uint64, c_uint64, P]),
    "kpms_ar_loglik_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "kpms_ar_loglik": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, P, P, c_size_t, c_int, P]),
    "kpms_ar_loglik_static": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, P, c_int, P]),
    "kpms_hmm_ffbs_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "kpms_hmm_ffbs": (c_int, [P, P, c_int64, P, P, c_int, c_int, c_int, c_int, c_int, P, P, P, P, c_size_t, P]),
    "kpms_resample_stateseqs_workspace_bytes": (c_size_t, [c_int] * 6),
    "kpms_resample_stateseqs": (c_int, [P, P, P, P, P, P] + [c_int] * 7 + [P, P, P, P, c_size_t, c_int, P]),
    "kpms_resample_latents_workspace_bytes": (c_size_t, [c_int] * 8),
    "kpms_resample_latents": (c_int, [P] * 12 + [c_int] * 9 + [P, P, P, P, c_size_t, P]),
    "kpms_ar_stats_workspace_bytes": (c_size_t, [c_int] * 5),
    "kpms_ar_stats": (c_int, [P, P, P, c_int, c_int, c_int, c_int, c_int, P, P, c_size_t, P]),
    "kpms_count_transitions": (c_int, [P, P, c_int64, c_int, c_int, c_int, P, P]),
    "kpms_crp_tablecounts": (c_int, [P, P, P, c_double, c_double, P, c_int64, P, c_int, c_int, P, P, P, P]),
    "kpms_crp_tablecounts_philox": (c_int, [P, P, P, c_double, c_double, c_uint64, c_uint64, c_uint64, c_int,
                                            P, P, P]),
    "kpms_obs_variance_stats_workspace_bytes": (c_size_t, [c_int] * 3),
    "kpms_obs_variance_stats": (c_int, [P] * 7 + [c_int] * 5 + [P, P, P, c_size_t, P]),
    "kpms_obs_params_stats_workspace_bytes": (c_size_t, [c_int] * 5),
    "kpms_log_joint_workspace_bytes": (c_size_t, []),
    "kpms_log_joint": (c_int, [P] * 13 + [c_double] * 3 + [c_int] * 7 + [P, P, c_size_t, P]),
    "kpms_obs_params_stats": (c_int, [P] * 7 + [c_int] * 5 + [P, P, P, c_size_t, P]),
    "kpms_resample_scales": (c_int, [P] * 6 + [c_double, c_double, P] + [c_int] * 5 + [P, P]),
    "kpms_resample_scales_stats": (c_int, [P] * 7 + [c_double, c_double, P, c_uint64, c_uint64, c_uint64] + [c_int] * 5 +
                                   [P, P, P, P, c_size_t, P]),
    "kpms_resample_heading_philox": (c_int, [P] * 6 + [c_uint64, c_uint64, c_uint64] + [c_int] * 5 + [P, P, P, P, P]),
    "kpms_resample_heading": (c_int, [P] * 7 + [c_int] * 6 + [P, P, P, P]),
    "kpms_resample_location_workspace_bytes": (c_size_t, [c_int] * 3),
    "kpms_resample_location": (c_int, [P] * 7 + [c_double, P, P] + [c_int] * 5 + [P, P, P, c_size_t, P]),
}

_lib = None


class KpmsError(RuntimeError):
    pass


def load():
    """Load the shared library (once) and attach prototypes.  Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KpmsError(
            "libkpms_b200.so not found at %s; run `python -m keypointmoseq_b200.build` "
            "(there is no CPU or PyTorch fallback for the CUDA path)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def set_option(name, value):
    """Debugging / cross-check switch of the library (see include/kpms_b200.h: kpms_set_option)."""
    check(load().kpms_set_option(name.encode(), int(value)), "kpms_set_option(%s)" % name)


def get_option(name):
    v = c_int(0)
    check(load().kpms_get_option(name.encode(), ctypes.byref(v)), "kpms_get_option(%s)" % name)
    return v.value


def check(status, what):
    if status != 0:
        msg = load().kpms_status_string(int(status)).decode()
        raise KpmsError("%s failed with status %d: %s" % (what, status, msg))


def ptr(t):
    """Raw device pointer of a contiguous CUDA tensor (or NULL for None)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise KpmsError("expected a CUDA tensor; the keypointmoseq_b200 path has no CPU fallback")
    if not t.is_contiguous():
        raise KpmsError("expected a contiguous tensor")
    return c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Workspace:
    """Grow-only per-device scratch buffers, keyed by name.  The library itself never
    allocates; this is the caller-side allocator the C-ABI asks for."""

    def __init__(self):
        self._bufs = {}

    def get(self, name, nbytes, device):
        key = (name, str(device))
        buf = self._bufs.get(key)
        if buf is None or buf.numel() < nbytes:
            self._bufs[key] = buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        return buf

    def clear(self):
        self._bufs.clear()


WORKSPACE = Workspace()
