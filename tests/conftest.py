import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "smallpt-enoki-optix_b200")
for p in (ROOT, PKG, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def built():
    """Make sure the oracle, the test-only host library and libwpt.so exist (built in-tree)."""
    need = [os.path.join(ROOT, "oracle", "libwpt_oracle.so"), os.path.join(PKG, "libwpt.so"),
            os.path.join(PKG, "libwpt_hosttest.so")]
    if not all(os.path.exists(p) for p in need):
        sys.path.insert(0, ROOT)
        import __graft_entry__
        __graft_entry__.build()
    return True


@pytest.fixture(scope="session")
def hosttest():
    import ctypes as C
    # WPT_HOSTTEST_LIB: tools/sanitize.sh points this at the ASan/UBSan build
    return C.CDLL(os.environ.get("WPT_HOSTTEST_LIB", os.path.join(PKG, "libwpt_hosttest.so")))


def pytest_sessionfinish(session, exitstatus):
    """WPT_LIB=.../libwpt_checked.so python -m pytest tests -m gpu: the whole GPU suite on the self-checking build —
    at the end every invariant counter of the traversal kernels must still be zero (tests/test_gpu_checked.py does this
    for tools/sanitize_target.py; this hook extends it to everything the suite launched in this process)."""
    if "checked" not in os.environ.get("WPT_LIB", ""):
        return
    try:
        import ctypes as C
        import wpt
        out = (C.c_uint32 * 8)()
        rc = wpt.lib().wpt_debug_check_failures(out, C.c_int(0))
        print("\nWPT_CHECKS counters after the session:", "unavailable" if rc else list(out))
        if rc == 0 and any(out):
            session.exitstatus = 1
    except Exception as e:          # no device / library: nothing to report
        print("\nWPT_CHECKS counters: not read (%s)" % e)
