import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HERE = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, HERE):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # GPU tests never silently skip on a GPU box; on a CPU box `-m "not gpu"` deselects them.
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture
def options():
    """Set library options (kpms_set_option) for one test; the previous values come back afterwards."""
    from keypointmoseq_b200 import _lib
    saved = {}

    def set_(name, value):
        if name not in saved:
            saved[name] = _lib.get_option(name)
        _lib.set_option(name, value)

    yield set_
    for name, value in saved.items():
        _lib.set_option(name, value)


@pytest.fixture(autouse=True)
def _fresh_warmup_adaptation(request):
    """The adaptive warm-up multipliers of the chunked passes are process state that follows the models a fit
    has seen; every GPU test starts from the defaults, so results do not depend on the test order."""
    if "gpu" not in request.keywords:
        yield
        return
    from keypointmoseq_b200 import gibbs as G
    G._HMM_ADAPT.clear()
    G._HMM_ADAPT.update(boost=1, streak=0, need=8, probing=False)
    G._LAT_ADAPT.clear()
    G._LAT_ADAPT.update(boost=1, streak=0, need=3, probing=False)
    yield
