import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "domain-mcp-server_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _ensure_built():
    """Built artefacts are git-ignored: a fresh checkout compiles them on first use (nvcc cross-compiles
    sm_100a without a GPU)."""
    import subprocess
    pkg = os.path.join(ROOT, "domain-mcp-server_b200")
    if not os.path.exists(os.path.join(pkg, "libdmgraph.so")):
        subprocess.check_call(["make", "-s", "-C", pkg, "libdmgraph.so"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


def pytest_configure(config):
    _ensure_built()
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import dmgraph
        return dmgraph.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
