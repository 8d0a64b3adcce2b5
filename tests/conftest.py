import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# The reference's third-party client is not installable here; tests run against the stand-in.
from oracle import ensure_kubernetes_importable  # noqa: E402

ensure_kubernetes_importable()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "needs_reference: imports /root/reference (build container only)")


def pytest_collection_modifyitems(config, items):
    from oracle.reference_loader import reference_available
    if reference_available():
        return
    skip = pytest.mark.skip(reason="/root/reference is not present on this machine")
    for it in items:
        if "needs_reference" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def native_pieces_built():
    """Build what is missing (nvcc and gcc are in the image on both machines); never silently skip."""
    from k8s_gpu_node_checker_b200 import native, pnl
    if not os.path.exists(native.LIB_PATH):
        native.build()
    if pnl._fastpack() is None:
        native.build_fastpack()


@pytest.fixture(scope="session")
def golden():
    import json
    gold = os.path.join(ROOT, "tests", "golden")

    def load(name):
        with open(os.path.join(gold, name)) as f:
            return json.load(f)
    return load
