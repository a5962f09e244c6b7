import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(autouse=True)
def _gn_ticket_counters_clean(request):
    """After every GPU test the GroupNorm ticket counters of libb200ad must be back at zero (a residue would make
    later GroupNorm launches finalize early on stale partial sums)."""
    yield
    if "gpu" not in request.keywords or os.environ.get("B200AD_NO_COUNTER_CHECK"):
        return
    import torch
    if not torch.cuda.is_available():
        return
    import b200ad  # noqa: F401
    from b200ad import lib
    nz = lib.load().b200ad_gn_counters_nonzero()
    assert nz <= 0, f"{nz} GroupNorm ticket counters left non-zero by {request.node.nodeid}"
