"""mgb_vector_dot_no_alloc (include/mgb200.h; no counterpart in the reference):
the dot product whose reduction scratch is a static __device__ array instead of cudaMalloc'ed buffers, for callers that
cannot allocate.  Called through the C ABI and compared with the existing entry point mgb_vector_dot on the same device
vectors, on two calls in a row (the calls share the scratch array, which keeps the previous call's contents), within the
tolerance of the other parity tests (1e-12 relative).  This file is named so that pytest collects it LAST:
a failure here must not hide the tests before it."""
import numpy as np
import pytest

from util import PARITY_RTOL, seeded

pytestmark = pytest.mark.gpu


def _ptr(v):
    """raw device address of a vector's storage (0 for an empty vector)"""
    return v.data_ptr if v.n_owned else 0


def _dot_pair(mg, n, seed):
    """(no-alloc result, existing entry point's result, numpy's) for seeded vectors of length n"""
    a, b = seeded(n, seed), seeded(n, seed + 100)
    x, y, r = mg.Vector(n), mg.Vector(n), mg.Vector(1)
    if n:
        x.from_numpy(a), y.from_numpy(b)
    r.from_numpy(np.array([123.456]))                       # stale value: the call must overwrite it
    mg.dot_no_alloc(_ptr(x), _ptr(y), n, _ptr(r))
    mg.synchronize()
    got = float(r.to_numpy()[0])
    want = x.dot(y) if n else 0.0                           # mgb_vector_dot
    return got, want, float(a @ b)


def test_no_alloc_dot_matches_the_existing_entry_point_on_two_calls_in_a_row(gpu):
    mg = gpu
    n = 1_000_003                                           # 245 blocks, ragged last block
    a, b = seeded(n, 1), seeded(n, 2)
    x, y, r = mg.Vector(n), mg.Vector(n), mg.Vector(1)
    x.from_numpy(a), y.from_numpy(b)
    want = x.dot(y)                                         # existing entry point (mgb_vector_dot)
    scale = float(np.linalg.norm(a) * np.linalg.norm(b))
    results = []
    for _ in range(2):                                      # two calls in a row on the shared scratch array
        r.from_numpy(np.array([-7.0]))
        mg.dot_no_alloc(_ptr(x), _ptr(y), n, _ptr(r))
        mg.synchronize()
        results.append(float(r.to_numpy()[0]))
    for got in results:
        assert abs(got - want) <= PARITY_RTOL * scale
        assert abs(got - float(a @ b)) <= PARITY_RTOL * scale
    assert results[0] == results[1]                         # fixed summation order: bitwise reproducible
    # a second, different pair straight afterwards: nothing of the first call's partial sums leaks into it
    y.from_numpy(a)
    mg.dot_no_alloc(_ptr(x), _ptr(y), n, _ptr(r))
    mg.synchronize()
    assert abs(float(r.to_numpy()[0]) - x.dot(y)) <= PARITY_RTOL * float(a @ a)


@pytest.mark.parametrize("n", [0, 1, 255, 4096, 4097, 65_537, 5_000_000])
def test_no_alloc_dot_sizes(gpu, n):
    got, want, ref = _dot_pair(gpu, n, 7 + n % 13)
    scale = max(1.0, n / 3.0)                               # |x||y| for uniform(-1,1) entries
    assert abs(got - want) <= PARITY_RTOL * scale and abs(got - ref) <= PARITY_RTOL * scale


def test_no_alloc_dot_on_a_caller_stream(gpu):
    import torch
    mg = gpu
    n = 300_000
    a, b = seeded(n, 3), seeded(n, 4)
    x, y, r = mg.Vector(n), mg.Vector(n), mg.Vector(1)
    x.from_numpy(a), y.from_numpy(b)
    mg.synchronize()
    s = torch.cuda.Stream()
    mg.dot_no_alloc(_ptr(x), _ptr(y), n, _ptr(r), s.cuda_stream)
    s.synchronize()
    assert abs(float(r.to_numpy()[0]) - x.dot(y)) <= PARITY_RTOL * float(np.linalg.norm(a) * np.linalg.norm(b))


def test_no_alloc_dot_rejects_sizes_beyond_the_static_scratch(gpu):
    mg = gpu
    x, r = mg.Vector(16), mg.Vector(1)
    with pytest.raises(mg.MgbError, match="exceeds the maximum"):
        mg.dot_no_alloc(_ptr(x), _ptr(x), mg.NO_ALLOC_MAX_N + 1, _ptr(r))
    with pytest.raises(mg.MgbError, match="null pointer"):
        mg.dot_no_alloc(_ptr(x), _ptr(x), 16, 0)
    # the limit itself is accepted by the size check (not run: 2 x 128 MiB vectors are not needed to prove the bound)
