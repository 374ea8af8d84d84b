"""f110_edt_build_noalloc (include/f110_b200.h): the EDT build for callers that cannot call cudaMalloc.  No counterpart in the
reference.  Its scratch is a fixed static device array inside the library, shared
by every call, so the test makes calls IN A ROW on maps of different sizes (the second call finds the first one's contents in
the scratch) and compares each output with the existing, allocating entry point f110_edt_build -- bit for bit, the tolerance
tests/test_gpu_edt.py uses for the EDT (exact integer squared distances, correctly rounded sqrt and multiply on both sides).
Collected last on purpose (file name)."""
import ctypes
import os

import numpy as np
import pytest

from conftest import MAPS

pytestmark = pytest.mark.gpu


def _build(fn, free_t, res, out_t, torch):
    from red_gym_b200 import _native as nat
    h, w = free_t.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream(free_t.device).cuda_stream)
    return fn(ctypes.c_void_p(free_t.data_ptr()), h, w, float(res), ctypes.c_void_p(out_t.data_ptr()), free_t.device.index, stream)


def test_edt_build_noalloc_matches_edt_build_on_calls_in_a_row():
    import torch
    from red_gym_b200 import _native as nat
    from red_gym_b200 import host
    L = nat.load()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(17)
    rnd = rng.random((700, 1301)) > 0.002          # sparse obstacles: long searches, a width that is no multiple of anything
    rnd[0, :] = False
    cases = [(host.MapData(os.path.join(MAPS, "example_map.yaml"), ".png", edt="gpu").free, 0.0625),      # 1600 x 1600, the bench map
             (rnd.astype(np.uint8), 0.05),
             (host.MapData(os.path.join(MAPS, "berlin.yaml"), ".png", edt="gpu").free, 0.05)]
    frees = [torch.from_numpy(np.ascontiguousarray(f)).to(dev) for f, _ in cases]
    want, got = [], []
    for f, (_, res) in zip(frees, cases):          # the existing entry point (allocates its scratch per call)
        o = torch.empty(f.shape, dtype=torch.float64, device=dev)
        nat.check(_build(L.f110_edt_build, f, res, o, torch))
        want.append(o)
        got.append(torch.full(f.shape, -1.0, dtype=torch.float64, device=dev))
    torch.cuda.synchronize()
    # calls in a row, no synchronisation between them (same stream: ordered), different sizes; then the first map again
    order = [0, 1, 2, 0]
    for k in order:
        got[k].fill_(-1.0)
        nat.check(_build(L.f110_edt_build_noalloc, frees[k], cases[k][1], got[k], torch))
    torch.cuda.synchronize()
    for k in range(len(cases)):
        assert torch.equal(got[k], want[k]), (k, float((got[k] - want[k]).abs().max()))
        assert float(got[k].min()) == 0.0 and float(got[k].max()) > 1.0
    # the call allocates nothing: device memory in use is the same before and after a further call
    free0, _ = torch.cuda.mem_get_info(dev)
    nat.check(_build(L.f110_edt_build_noalloc, frees[1], cases[1][1], got[1], torch))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0 and torch.equal(got[1], want[1])


def test_edt_build_noalloc_rejects_sizes_beyond_its_scratch():
    import torch
    from red_gym_b200 import _native as nat
    L = nat.load()
    dev = torch.device("cuda", 0)
    big = torch.ones((2049, 2048), dtype=torch.uint8, device=dev)           # one row more than F110_EDT_NOALLOC_MAX_CELLS
    out = torch.zeros((2049, 2048), dtype=torch.float64, device=dev)
    rc = _build(L.f110_edt_build_noalloc, big, 0.05, out, torch)
    assert rc == nat.ERR_INVALID and "exceed the static scratch" in nat.last_error() and "f110_edt_build" in nat.last_error()
    assert float(out.abs().max()) == 0.0                                    # nothing was launched
    wide = torch.ones((2, 16385), dtype=torch.uint8, device=dev)           # few cells, but wider than F110_EDT_NOALLOC_MAX_WIDTH
    rc = _build(L.f110_edt_build_noalloc, wide, 0.05, torch.zeros((2, 16385), dtype=torch.float64, device=dev), torch)
    assert rc == nat.ERR_INVALID
    ok = torch.ones((2048, 2048), dtype=torch.uint8, device=dev)           # exactly the maximum works
    ok[5, 7] = 0
    a = torch.empty((2048, 2048), dtype=torch.float64, device=dev)
    b = torch.empty_like(a)
    nat.check(_build(L.f110_edt_build_noalloc, ok, 0.05, a, torch))
    nat.check(_build(L.f110_edt_build, ok, 0.05, b, torch))
    assert torch.equal(a, b) and float(a[5, 7]) == 0.0
