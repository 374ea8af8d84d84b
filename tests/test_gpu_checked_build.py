"""Memory-safety evidence without compute-sanitizer (closed on the B200 pool): the tests that exercise the code whose speed
comes from removed or hoisted bounds tests -- the padded fp32 EDT addressed through a biased base pointer and the mantissa
floor, the fp64 path's hoisted map test, the trig / direction table indices, the opponent-cast queue, the sort bins, beam and
noise-row indices -- are run once more in a subprocess against libf110_b200_checked.so, the same sources compiled with
-DF110_CHECKED, in which every such address / index is tested on the device (csrc/f110_dev.h: F110_CHECK).  Zero violations
and the same green tests; the deliberate self-test violation proves the checks are compiled in."""
import os
import subprocess
import sys

import pytest

from conftest import REPO

pytestmark = pytest.mark.gpu


def test_bounds_checked_build_runs_clean():
    lib = os.path.join(REPO, "red_gym_b200", "libf110_b200_checked.so")
    assert os.path.exists(lib), "build it with `python red_gym_b200/build.py --checked` (build() of __graft_entry__ does)"
    env = dict(os.environ)
    env.pop("F110_B200_LIB", None)
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", "checked_run.py")], capture_output=True, text=True, env=env, timeout=1500)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    line = [l for l in r.stdout.splitlines() if l.startswith("CHECKED RUN:")][-1]
    print(line)
    assert "'selftest': 1" in line and "clean" in line


def test_normal_build_has_no_checks():
    import ctypes
    from red_gym_b200 import _native as nat
    if os.environ.get("F110_B200_LIB"):
        pytest.skip("another build of the library is selected (tools/checked_run.py runs the suite against the checked one)")
    L = nat.load()
    assert L.f110_build_is_checked() == 0
    out = (ctypes.c_uint64 * 8)()
    nat.check(L.f110_debug_violations(0, out, 1))
    assert list(out) == [0] * 8          # the self-test kernel counts nothing here
