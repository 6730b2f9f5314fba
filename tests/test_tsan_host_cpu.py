"""ThreadSanitizer run of the host mirror's GraphService (csrc/host_graph.cc; GS:55-136): writers put /
reload graphs while readers get and release them.  The device calls are stand-ins (tests/native/
tsan_service.cc), so this runs on CPU; it fails on any data race TSAN reports, on a leaked or torn graph
handle, and when the toolchain has no TSAN runtime it is skipped."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_graph_service_is_race_free_under_tsan(tmp_path):
    exe = str(tmp_path / "tsan_service")
    cmd = ["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-fPIE", "-pie",
           os.path.join(ROOT, "tests", "native", "tsan_service.cc"),
           os.path.join(ROOT, "domain-mcp-server_b200", "csrc", "host_graph.cc"), "-o", exe, "-lpthread"]
    b = subprocess.run(cmd, capture_output=True, text=True)
    if b.returncode != 0 and ("tsan" in b.stderr.lower() or "sanitize" in b.stderr.lower()):
        pytest.skip("no ThreadSanitizer runtime: " + b.stderr[-200:])
    assert b.returncode == 0, b.stderr[-2000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66"))
    assert "WARNING: ThreadSanitizer" not in r.stderr, r.stderr[-4000:]
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr[-2000:])
    assert "bad=0" in r.stdout and "live_after_free=0" in r.stdout
