"""-m gpu: the self-checking build.  compute-sanitizer is closed on this GPU pool (its own message: "closed on this pool
and stays closed"), so libwpt_checked.so — the same sources with -DWPT_CHECKS — makes the traversal kernels count
violations of the invariants memcheck/racecheck would watch in the warp-shared queue code: queue / item / stack slots in
range, every block index inside the tree, every fetched ray index inside the queue, the queue empty after a drain, the
scanned item total equal to the triangles expanded.  All workloads of tools/sanitize_target.py (every kernel of the
pipeline, both flush schemes, the local-memory stack overflow, masked per-ray-window intersect, a full-size C3 frame with
and without the fused tail) must leave every counter at zero — and produce the product build's numbers."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(lib):
    env = dict(os.environ)
    if lib:
        env["WPT_LIB"] = os.path.join(ROOT, "smallpt-enoki-optix_b200", lib)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_target.py"), "all"], env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    return [ln for ln in r.stdout.splitlines() if ln.strip()]


def test_checked_build_reports_no_violation_and_the_same_results():
    checked = _run("libwpt_checked.so")
    assert checked[-1] == "checks [0, 0, 0, 0, 0, 0, 0, 0]", checked[-1]
    product = _run(None)
    assert product[-1] == "checks unavailable"          # the product build carries no checks
    assert checked[:-1] == product[:-1]                  # same segment counts and film sums, line by line
    assert len(checked) >= 10
