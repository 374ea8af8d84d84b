"""Installs the UNMODIFIED reference where the GPU box can import it: oracle/_ref/  (git-ignored build output).

    python oracle/install_ref.py          # called by __graft_entry__.build() when /root/reference is present

`/root/reference` does not exist on the GPU box, but `oracle/_ref/` travels there with the snapshot (like the built
.so files), so `bench.py --impl reference` and bench.py's `cpu_baseline` leg can time the reference's own
Numba/NumPy path (kind "reference") instead of the C restatement (kind "port").  Test infrastructure only: nothing
under red_gym_b200/ or f110_gym/ reads this directory.

Recipe (the contract's): `pip install --no-index --no-build-isolation --no-deps --target oracle/_ref <copy of
/root/reference/gym>` -- from a copy because the reference tree is read-only and setuptools writes build/ and egg-info
next to setup.py; --no-deps because the pinned requirements (gym==0.19.0, numpy<=1.22, pyglet==1.5.20, pyopengl) are not
in the offline wheelhouse: numpy / numba / scipy / Pillow / pyyaml of this image are used instead (no failures
from the version skew), and `gym` / `pyglet`, which the reference imports but never uses on this path, are the inert
stand-ins of oracle/ref_shim.py, exactly as for the golden generation.
"""
import json
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SRC = os.environ.get("F110_REFERENCE_ROOT", "/root/reference")


def install(force=False):
    if not os.path.isfile(os.path.join(SRC, "gym", "setup.py")):
        return None                       # GPU box: use what travelled with the snapshot
    marker = os.path.join(DEST, "INSTALL.json")
    if os.path.exists(marker) and not force:
        return DEST
    shutil.rmtree(DEST, ignore_errors=True)
    with tempfile.TemporaryDirectory() as tmp:
        shutil.copytree(os.path.join(SRC, "gym"), os.path.join(tmp, "gym"))
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--quiet", "--no-compile",
               "--target", DEST, os.path.join(tmp, "gym")]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0 or not os.path.isfile(os.path.join(DEST, "f110_gym", "envs", "f110_env.py")):
        raise RuntimeError("pip install of the reference failed (rc %d):\n%s" % (r.returncode, r.stdout))
    with open(marker, "w") as f:
        json.dump({"source": os.path.join(SRC, "gym"), "command": " ".join(cmd[:-1] + ["<copy of %s/gym>" % SRC]),
                   "installed": sorted(os.listdir(DEST))}, f, indent=1)
    return DEST


if __name__ == "__main__":
    print(install(force="--force" in sys.argv))
