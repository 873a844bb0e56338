"""Stand-in for the compute-sanitizer gate: the tool is closed on this GPU pool
(profiles/r02_sanitizer_closed_on_this_pool.log), so the reduced workload of tools/sanitize_case.py - every kernel family once, small
sizes, results compared with the oracle, one batch repeated ten times for bit-identical output - runs with 256-byte guard bands behind
every device buffer of the handles (ORBX_GUARD=1, orbx_debug_check_guards) and under the runtime switches that select the other code
paths: 256-thread strips, a tiny strip list (every strip overflows into the per-cell kernel), the per-cell FAST alone, ordinary-load
staging instead of TMA, the other describe kernels, the split quadtree pipeline, the matcher's CTA widths, stream priorities."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

ENVS = [{}, {"ORBX_STRIP_THREADS": "256"}, {"ORBX_STRIP_LIST": "256"}, {"ORBX_STRIP_MAXW": "235", "ORBX_STRIP_LIST": "4096", "ORBX_STRIP_MINB": "8"},
        {"ORBX_STRIP_MINB": "12", "ORBX_STRIP_MAXW": "75"}, {"ORBX_FAST_STRIP": "0"}, {"ORBX_FAST_TMA": "0"},
        {"ORBX_MATCH_R": "0"}, {"ORBX_BLUR_MODE": "3"},
        # round 2, later: the round-1 describe kernel and the bulk-copy form of the pipelined one, the split quadtree pipeline (on its own and
        # with every strip overflowing, i.e. the gated late pass), narrow CTAs in the consecutive matcher, the main stream above the auxiliary
        {"ORBX_DESC_MODE": "0"}, {"ORBX_DESC_MODE": "2"}, {"ORBX_PIPE": "3"}, {"ORBX_PIPE": "3", "ORBX_STRIP_LIST": "256"},
        {"ORBX_MATCH_NT": "64"}, {"ORBX_PRIO": "1"}, {"ORBX_STEREO_ROWS": "0"}]


@pytest.mark.parametrize("extra", ENVS, ids=lambda e: "+".join(f"{k}={v}" for k, v in e.items()) or "default")
def test_reduced_workload_with_guard_bands(extra):
    env = dict(os.environ, ORBX_GUARD="1", **extra)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_case.py")], capture_output=True, text=True, timeout=600,
                       env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "sanitize_case ok" in r.stdout and "bit-identical" in r.stdout
