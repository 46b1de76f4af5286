"""wf2cu_sum_all_noalloc — the allocation-free `sum :axis t` (scratch in a file-scope device array, include/waffe2cuda.h)
— against the existing entry point wf2cu_reduce_sum, through the C ABI.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

# the bar of the reduction parity tests (tests/test_gpu_parity.py): 1e-5 of sum|x| against an fp64 sum
REL = 1e-5


def _existing(gpu, x, n, accumulate=0, divisor=0.0, start=0.0):
    lib = gpu._lib.load()
    out = gpu.from_numpy(np.full((1,), start, np.float32))
    xv = gpu._lib.make_view(x.vec.ptr, 0, [n], [1])
    ov = gpu._lib.make_view(out.vec.ptr, 0, [n], [0])
    gpu._lib.check(lib.wf2cu_reduce_sum(C.byref(xv), C.byref(ov), accumulate, divisor))
    return float(out.to_numpy()[0])


def _noalloc(gpu, x, n, accumulate=0, divisor=0.0, start=0.0, x_off=0):
    lib = gpu._lib.load()
    out = gpu.from_numpy(np.full((1,), start, np.float32))
    gpu._lib.check(lib.wf2cu_sum_all_noalloc(x.vec.ptr, x_off, n, out.vec.ptr, 0, accumulate, divisor))
    return float(out.to_numpy()[0])


@pytest.mark.parametrize("n", [4096 * 4096, (1 << 20) + 12345, 1 << 18, 5000, 3])
def test_noalloc_sum_matches_existing_entry_point_twice(gpu, n):
    rng = np.random.default_rng(n % 1000)
    for call in range(2):                                   # two calls in a row share the static scratch
        X = rng.standard_normal(n).astype(np.float32)
        x = gpu.from_numpy(X)
        truth = float(X.astype(np.float64).sum())
        bar = REL * float(np.abs(X).astype(np.float64).sum())
        want = _existing(gpu, x, n)
        got = _noalloc(gpu, x, n)
        assert abs(got - want) <= bar, (call, got, want)
        assert abs(got - truth) <= bar, (call, got, truth)
        if n >= 1 << 18:                                     # both run the flat kernel on the same grid: same bits
            assert got == want, (call, got, want)


def test_noalloc_accumulate_divisor_and_offset(gpu):
    n = (1 << 19) + 4
    rng = np.random.default_rng(5)
    X = rng.standard_normal(n + 8).astype(np.float32)
    x = gpu.from_numpy(X)
    for _ in range(2):
        want = (np.float32(2.5) + np.float32(X[8:].astype(np.float64).sum())) / np.float32(n)
        got = _noalloc(gpu, x, n, accumulate=1, divisor=float(n), start=2.5, x_off=8)
        assert abs(got - float(want)) <= REL * float(np.abs(X).sum()) / n + 1e-6


def test_noalloc_refuses_beyond_maximum_and_misaligned(gpu):
    lib = gpu._lib.load()
    x = gpu.from_numpy(np.ones(64, np.float32))
    out = gpu.from_numpy(np.zeros(1, np.float32))
    rc = lib.wf2cu_sum_all_noalloc(x.vec.ptr, 0, (1 << 28) + 1, out.vec.ptr, 0, 0, 0.0)
    assert rc != 0
    with pytest.raises(gpu._lib.Wf2CudaError, match="maximum"):
        gpu._lib.check(rc)
    with pytest.raises(gpu._lib.Wf2CudaError, match="aligned"):
        gpu._lib.check(lib.wf2cu_sum_all_noalloc(x.vec.ptr, 1, 8, out.vec.ptr, 0, 0, 0.0))
    # and a good call still works afterwards
    assert _noalloc(gpu, x, 64) == 64.0
