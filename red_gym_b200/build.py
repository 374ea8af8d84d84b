"""Builds libf110_b200.so in-tree with nvcc for sm_100a (B200).  `python red_gym_b200/build.py`.

Three translation units: the fp64 kernels are compiled with -fmad=false (the reference's Numba
code has no FMA contraction and the parity path must round the same way); the fp32 fast path
keeps FMA; the C-ABI glue owns memory and launch order.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libf110_b200.so")
LIB_CHECKED = os.path.join(HERE, "libf110_b200_checked.so")      # -DF110_CHECKED: the same code with its own bounds checks (f110_dev.h)
LIB_PHASES = os.path.join(HERE, "libf110_b200_phases.so")        # -DF110_PHASES: %globaltimer stamps inside k_step_fused64 (tools/phase_probe.py)
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
UNITS = [
    ("f110_kernels_fp64.cu", ["-fmad=false"]),
    ("f110_kernels_fp32.cu", []),
    ("f110_order.cu", []),
    ("f110_bitmap.cu", ["-fmad=false"]),
    ("f110_pursuit.cu", ["-fmad=false"]),
    ("f110_edt.cu", ["-fmad=false"]),
    ("f110_render.cu", ["-fmad=false"]),
    ("f110_probe.cu", []),
    ("f110_abi.cu", []),
]


def nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    out = [os.path.join(CSRC, u) for u, _ in UNITS]
    out += [os.path.join(CSRC, "f110_dev.h"), os.path.join(CSRC, "f110_bitmap_dev.h"), os.path.join(CSRC, "f110_pursuit_dev.h"), os.path.join(CSRC, "f110_render_dev.h"), os.path.join(CSRC, "f110_scan32_dev.h"), os.path.join(os.path.dirname(HERE), "include", "f110_b200.h")]
    return out


def stale(lib=LIB):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(s) > t for s in sources())


def build(force=False, verbose=False, checked=False, phases=False):
    LIB = LIB_PHASES if phases else LIB_CHECKED if checked else globals()["LIB"]
    if not force and not stale(LIB):
        return LIB
    cc = nvcc()
    objdir = os.path.join(HERE, "build_phases" if phases else "build_checked" if checked else "build")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for unit, extra in UNITS:
        obj = os.path.join(objdir, unit.replace(".cu", ".o"))
        cmd = [cc] + ARCH + COMMON + extra + (["-DF110_PHASES"] if phases else ["-DF110_CHECKED"] if checked else []) + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(CSRC, unit), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose and out:
            print(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), out))
    cmd = [cc] + ARCH + ["-shared", "-o", LIB] + objs + ["-cudart", "static"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, checked="--checked" in sys.argv, phases="--phases" in sys.argv))
