import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _name_the_test_in_the_parity_report(request):
    import helpers
    helpers._CURRENT_TEST[0] = request.node.nodeid
    yield


def pytest_sessionfinish(session, exitstatus):
    """Observed error statistics of every float comparison (tests/helpers.compare_float) -> gpurun_out/."""
    import json
    try:
        import helpers
    except Exception:
        return
    out = os.path.join(ROOT, "gpurun_out")
    if helpers.REPORT and os.path.isdir(out):
        import torch
        tag = "gpu" if torch.cuda.is_available() else "cpu"
        with open(os.path.join(out, f"parity_report_{tag}.json"), "w") as f:
            json.dump(helpers.REPORT, f, indent=1, sort_keys=True)
