"""CPU-side checks of the C-ABI boundary: the shared library loads, exports every symbol
that include/kpms_b200.h declares (and nothing the binding expects is missing), and the
argument validation answers with the documented status codes — no kernel is launched."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from keypointmoseq_b200 import build, _lib
    build.build()
    return _lib.load()


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "kpms_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(kpms_[a-z0-9_]+)\s*\(", txt)))


def test_header_and_binding_agree(lib):
    from keypointmoseq_b200 import _lib
    syms = _header_symbols()
    assert len(syms) >= 20
    assert sorted(_lib.SIGNATURES) == syms
    for s in syms:
        assert hasattr(lib, s), s


def test_every_entry_point_cites_the_reference():
    txt = open(os.path.join(ROOT, "include", "kpms_b200.h")).read()
    assert txt.count("ref:") >= 10


def test_status_strings_and_version(lib):
    assert lib.kpms_version() >= 100
    assert lib.kpms_status_string(0) == b"ok"
    for code in (-1, -2, -3, -4, -5):
        assert lib.kpms_status_string(code).startswith(b"kpms:")


def test_argument_validation_without_gpu(lib):
    null = ctypes.c_void_p(None)
    one = ctypes.c_void_p(256)
    assert lib.kpms_ar_loglik(null, null, null, 1, 10, 10, 3, 5, null, null, 0, 0, null) == -1
    assert lib.kpms_ar_loglik(one, one, one, 1, 3, 10, 3, 5, one, one, 1 << 20, 0, null) == -2   # T <= L
    assert lib.kpms_ar_loglik(one, one, one, 1, 10, 9, 3, 5, one, one, 1 << 20, 0, null) == -3   # D,L not built
    assert lib.kpms_ar_loglik(one, one, one, 1, 10, 10, 3, 5, one, one, 8, 0, null) == -4        # workspace
    assert lib.kpms_hmm_ffbs(one, one, 5, one, one, 1, 10, 4, 0, 0, one, one, null, one, 1 << 20, null) == -2  # stride < Tm
    assert lib.kpms_hmm_ffbs(one, one, 10, one, one, 1, 10, 2000, 0, 0, one, one, null, one, 1 << 30, null) == -3
    assert lib.kpms_fill_gamma(one, 10, 0.5, 1, 1, null) == -3
    assert lib.kpms_resample_latents_workspace_bytes(0, 10, 8, 2, 10, 3, 5, 0) == 0


def test_workspace_queries_scale(lib):
    a = lib.kpms_resample_latents_workspace_bytes(10, 1000, 8, 2, 10, 3, 50, 0)
    b = lib.kpms_resample_latents_workspace_bytes(20, 1000, 8, 2, 10, 3, 50, 0)
    assert 1.9 < b / a < 2.1 and a > 10 * 998 * 495 * 8
    assert lib.kpms_hmm_ffbs_workspace_bytes(10, 100, 50, 0) >= 10 * 100 * 50 * 4
    assert lib.kpms_ar_stats_workspace_bytes(10, 100, 10, 3, 50) % 256 == 0


def test_product_has_no_oracle_import():
    pkg = os.path.join(ROOT, "keypointmoseq_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|kpms_oracle|importlib.*oracle", src, re.M), f
