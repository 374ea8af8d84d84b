"""A short run of each configuration fuzzer (tools/config_fuzz.py, step_fuzz.py, fp32_fuzz.py; the long runs are recorded in
profiles/r2h_config_fuzz.txt): random beam counts, fields of view, maps, origins, agent counts, integrators, time steps and
vehicle parameters against the oracle (fp64: bit-exact scans, exact flags) and fp32 against fp64 -- in a subprocess each, against
the bounds-checked build, which must count no violation."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECKED = os.path.join(REPO, "red_gym_b200", "libf110_b200_checked.so")


@pytest.mark.parametrize("tool,configs,seed", [("config_fuzz.py", 30, 101), ("step_fuzz.py", 30, 102), ("fp32_fuzz.py", 12, 103)])
def test_fuzzer_finds_nothing(tool, configs, seed):
    if not os.path.exists(CHECKED):
        pytest.skip("libf110_b200_checked.so not built (python red_gym_b200/build.py --checked)")
    env = dict(os.environ, F110_B200_LIB=CHECKED)
    r = subprocess.run([sys.executable, os.path.join(REPO, "tools", tool), str(configs), str(seed)], capture_output=True, text=True, env=env, timeout=900)
    tail = "\n".join(r.stdout.splitlines()[-6:]) + r.stderr[-2000:]
    assert r.returncode == 0, tail
    last = r.stdout.splitlines()[-1]
    assert last.startswith("checked build: True") and "configs with a finding: 0 of %d" % configs in last, last
