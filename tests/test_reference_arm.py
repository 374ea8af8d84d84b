"""The CPU arm of bench.py (`--impl reference`, `cpu_baseline`): the unmodified reference F110Env, one process per core, imported
from /root/reference/gym (build container) or oracle/_ref (GPU box).  Measurement infrastructure; runs on the CPU."""
import json
import os
import subprocess
import sys

import pytest

from conftest import REPO

sys.path.insert(0, os.path.join(REPO, "oracle"))
import ref_runner  # noqa: E402

ok, why = ref_runner.available()
pytestmark = pytest.mark.skipif(not ok, reason="reference not importable here: %s" % why)


def test_reference_runner_times_the_unmodified_env():
    r = ref_runner.run(agents=1, seconds=1.0, solo_seconds=0.5, procs=2)
    assert r["kind"] == "reference" and r["cores"] == 2 and r["unit"] == "agent-steps/s"
    assert r["value"] > 100.0 and r["single_process_value"] > 50.0        # a Numba F110Env does thousands of steps per second
    assert r["sample_steps"] >= 2 * 16 and "unmodified reference" in r["sample"]
    r2 = ref_runner.run(agents=2, seconds=0.5, solo_seconds=0.0, procs=1)
    assert r2["single_process_value"] is None and r2["value"] > 20.0


def test_reference_arm_line_has_the_contract_keys_and_imports_no_product_code():
    env = dict(os.environ)
    env["F110_REF_PROCS"] = "2"
    code = ("import sys, json; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '3'];"
            "import runpy; runpy.run_path(%r, run_name='__main__');"
            "bad = [m for m in sys.modules if m.startswith('red_gym_b200')]; print('PRODUCT_MODULES', bad)" % os.path.join(REPO, "bench.py"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600, cwd=REPO)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "agent-steps/s" and line["higher_is_better"] is True
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["config"]["workload"] == "config3" and "cpu_model" in line["cpu_baseline"] and line["cpu_baseline"]["single_process_value"]
    assert "PRODUCT_MODULES []" in r.stdout              # the arm never imports the product (nor loads its .so)
