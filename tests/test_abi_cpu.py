"""CPU: the C-ABI library loads, exports every symbol include/dmgraph.h declares, and fails
loudly (no fallback) when there is no CUDA device."""
import ctypes as C
import os
import re

import pytest

import dmgraph

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "dmgraph.h")).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dmg_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    syms = header_symbols()
    assert len(syms) >= 29
    L = C.CDLL(dmgraph.LIB_PATH)
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing


def test_binding_covers_header():
    L = dmgraph.lib()
    for s in header_symbols():
        assert getattr(L, s).argtypes is not None, s


def test_no_cpu_fallback_without_device():
    if dmgraph.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(dmgraph.DmgError) as ei:
        dmgraph.Engine(0)
    assert ei.value.code == dmgraph.ENODEV
    assert "no CPU fallback" in str(ei.value)


def test_product_package_does_not_import_oracle():
    pkg = os.path.join(ROOT, "domain-mcp-server_b200")
    for base, _d, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                assert "oracle" not in open(os.path.join(base, f)).read().replace("SURVEY", ""), f
