"""Import shim for the UNMODIFIED reference (test infrastructure, container-only).

The reference (`/root/reference/gym/f110_gym`) imports `gym` (only for the
`gym.Env` base class and `register`, f110_env.py:28-30, __init__.py:1) and
`pyglet` (import-time options, f110_env.py:41-43); neither is installed.  This
module puts inert stand-ins into `sys.modules` and the reference tree on
`sys.path` so the reference hot path runs as-is.  Nothing here is product code;
it is used only by `oracle/gen_golden.py` and by tests that run in the build
container (the GPU box has no /root/reference).
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("F110_REFERENCE_ROOT", "/root/reference")
INSTALLED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")     # oracle/install_ref.py


def package_parent():
    """Directory to put on sys.path so that `import f110_gym` finds the reference: the reference tree when it is there
    (build container), else the unmodified copy oracle/install_ref.py placed under oracle/_ref (GPU box)."""
    for parent in (os.path.join(REFERENCE_ROOT, "gym"), INSTALLED):
        if os.path.isfile(os.path.join(parent, "f110_gym", "envs", "f110_env.py")):
            return parent
    return None


def available() -> bool:
    return package_parent() is not None


def install():
    """Make `import f110_gym` resolve to the reference package."""
    if not available():
        raise RuntimeError("reference not present at %s nor under %s" % (REFERENCE_ROOT, INSTALLED))
    os.environ.setdefault("NUMBA_CACHE_DIR", "/tmp/f110_ref_numba_cache")
    if "gym" not in sys.modules:
        gym = types.ModuleType("gym")

        class Env(object):
            metadata = {}

        gym.Env = Env

        def make(env_id, **kw):
            from f110_gym.envs import F110Env
            return F110Env(**kw)

        gym.make = make
        for sub in ("error", "spaces", "utils", "envs"):
            m = types.ModuleType("gym." + sub)
            setattr(gym, sub, m)
            sys.modules["gym." + sub] = m
        seeding = types.ModuleType("gym.utils.seeding")
        gym.utils.seeding = seeding
        sys.modules["gym.utils.seeding"] = seeding
        reg = types.ModuleType("gym.envs.registration")
        reg.register = lambda **kw: None
        gym.envs.registration = reg
        sys.modules["gym.envs.registration"] = reg
        sys.modules["gym"] = gym
    if "pyglet" not in sys.modules:
        pyglet = types.ModuleType("pyglet")
        pyglet.options = {}
        gl = types.ModuleType("pyglet.gl")
        gl.GL_POINTS = 0
        pyglet.gl = gl
        sys.modules["pyglet"] = pyglet
        sys.modules["pyglet.gl"] = gl
    ref_pkg = package_parent()
    # our own drop-in package is also called f110_gym: the reference must win here
    for name in [k for k in sys.modules if k == "f110_gym" or k.startswith("f110_gym.")]:
        del sys.modules[name]
    if ref_pkg in sys.path:
        sys.path.remove(ref_pkg)
    sys.path.insert(0, ref_pkg)
    import warnings
    warnings.filterwarnings("ignore", message="Chosen integrator is RK4")
    return sys.modules["gym"]
