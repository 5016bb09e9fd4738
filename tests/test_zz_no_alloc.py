"""sphb200_crossing_time_noalloc (the entry point that allocates no device memory) against the ordinary
sphb200_crossing_time, both through the C ABI, on two calls in a row.  The file sorts last on purpose:
a failure here must not hide another test."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-10     # the tolerance of the other floating-point parity tests (tests/test_gpu_parity.py)
MAX_ENTRIES = 4 << 20


def _inputs(n, k, seed, torch, dev):
    rng = np.random.default_rng(seed)
    nb = rng.integers(0, n, size=(n, k)).astype(np.int64)
    nb[:, 0] = np.arange(n)
    nb[::7, 3] = -1 - (np.arange(n)[::7] % n)              # numpy-style negative indices
    vel = rng.normal(0.0, 1.0e3, size=(n, 3))
    sizes = rng.uniform(1.0e14, 3.0e14, size=n)
    pt = (rng.uniform(size=n) < 0.2).astype(np.float64)    # 20 % dust: ignored by the minimum
    to = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return to(nb), to(vel), to(sizes), to(pt)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def test_crossing_time_noalloc_matches_the_ordinary_entry_point(built_lib):
    import torch
    from sphb200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ctx = _lib.Context(0)                                   # for the ordinary entry point only
    dt_0 = 3.0e9
    try:
        out = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
        bad = torch.full((2,), -1, dtype=torch.int32, device=dev)
        # two calls in a row with different inputs: the second must not see anything of the first in the scratch
        cases = [(20000, 40, 11), (3000, 17, 12)]
        want = []
        for c, (n, k, seed) in enumerate(cases):
            nb, vel, sizes, pt = _inputs(n, k, seed, torch, dev)
            free_before = torch.cuda.mem_get_info(dev)[0]
            _lib.check(lib.sphb200_crossing_time_noalloc(0, n, k, _ptr(nb), _ptr(vel), _ptr(sizes), _ptr(pt), dt_0,
                                                         ctypes.c_void_p(out.data_ptr() + 8 * c),
                                                         ctypes.c_void_p(bad.data_ptr() + 4 * c), None))
            torch.cuda.synchronize(dev)
            if c == 1:                                      # (the first call may load the module's kernels)
                assert torch.cuda.mem_get_info(dev)[0] >= free_before, "the call allocated device memory"
            ref = ctypes.c_double(0.0)
            _lib.check(lib.sphb200_crossing_time(ctx.handle, n, k, _ptr(nb), _ptr(vel), _ptr(sizes), _ptr(pt), dt_0,
                                                 ctypes.byref(ref), None))
            want.append(ref.value)
        got = out.cpu().numpy()
        assert bad.cpu().tolist() == [0, 0]
        for g, w in zip(got, want):
            assert w > 0.0 and abs(g - w) <= RTOL * abs(w), (g, w)
        assert want[0] != want[1]
        # no gas particle at all: dt_0 / 10 (nsc:786), and an out-of-range index is counted and skipped
        n, k = 500, 8
        nb, vel, sizes, pt = _inputs(n, k, 13, torch, dev)
        pt.fill_(1.0)
        nb[5, 2] = n + 3
        _lib.check(lib.sphb200_crossing_time_noalloc(0, n, k, _ptr(nb), _ptr(vel), _ptr(sizes), _ptr(pt), dt_0,
                                                     _ptr(out), _ptr(bad), None))
        torch.cuda.synchronize(dev)
        assert float(out[0]) == dt_0 / 10.0 and int(bad[0]) == 1
        # sizes beyond the fixed scratch are refused with a clear error, before anything is launched
        n_big = MAX_ENTRIES // 40 + 1
        rc = lib.sphb200_crossing_time_noalloc(0, n_big, 40, _ptr(nb), _ptr(vel), _ptr(sizes), _ptr(pt), dt_0,
                                               _ptr(out), _ptr(bad), None)
        assert rc != 0
        assert b"exceeds the fixed scratch" in lib.sphb200_last_error()
    finally:
        ctx.close()
