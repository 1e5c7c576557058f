"""halo_scalar_dot_noalloc_dev (include/halo_b200.h): the scalar_dot reduction for callers that cannot allocate.  Its
scratch is a fixed file-scope device array of the library; it must give exactly what the existing entry point
halo_scalar_dot gives (field arithmetic is exact, so the tolerance is zero, as in every other test here), on calls in a
row that share that array, and reject sizes beyond the array's capacity.

Collected last on purpose: a failure here hides no other test."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 1 << 24


@pytest.fixture(scope="module")
def hb():
    import halo_b200
    halo_b200.init()
    return halo_b200


def _dev(torch, a):
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint64).view(np.int64).copy()).cuda()
    torch.cuda.synchronize()
    return t


def _call(hb, torch, field, xs, ys, n, out):
    from halo_b200._lib import lib, check
    check(lib.halo_scalar_dot_noalloc_dev(ctypes.c_int(field), ctypes.c_void_p(xs.data_ptr() if xs is not None else 0),
                                          ctypes.c_void_p(ys.data_ptr() if ys is not None else 0), ctypes.c_size_t(n),
                                          ctypes.c_void_p(out.data_ptr())))
    check(lib.halo_device_synchronize())
    return out.cpu().numpy().view(np.uint64).copy()


@pytest.mark.parametrize("field", [0, 1])
def test_two_calls_in_a_row_equal_the_existing_entry_point(hb, oracle, field):
    import torch
    O = oracle
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    # a long call first (18 scratch slots written), then a short one (1 slot): the second must not see the first's
    # leftovers; then a ragged multi-slot size and the slot boundary itself
    for n, seed in ((70001, 3), (1000, 5), (3 * 4096 + 1, 7), (4096, 9), (1, 11)):
        a, b = O.random_scalars(field, n, seed), O.random_scalars(field, n, seed + 100)
        exp = hb.group.scalar_dot(field, a, b)                       # the existing entry point (host in, host out)
        da, db = _dev(torch, a), _dev(torch, b)
        got = _call(hb, torch, field, da, db, n, out)
        assert (got == exp).all(), n
        got2 = _call(hb, torch, field, da, db, n, out)               # the same call again on the shared scratch
        assert (got2 == exp).all(), n
    # ground truth for one of them with Python integers
    n = 1000
    a, b = O.random_scalars(field, n, 5), O.random_scalars(field, n, 105)
    want = sum(O.from_mont(field, x) * O.from_mont(field, y) for x, y in zip(a, b)) % O.MODULUS[field]
    assert O.from_mont(field, _call(hb, torch, field, _dev(torch, a), _dev(torch, b), n, out)) == want
    # n = 0: the sum is zero and the output is written
    out.fill_(-1)
    assert (_call(hb, torch, field, None, None, 0, out) == 0).all()


def test_allocates_nothing_and_rejects_sizes_beyond_the_scratch(hb, oracle):
    import torch
    from halo_b200._lib import lib, HALO_EINVAL
    O = oracle
    n = 1 << 16
    a, b = _dev(torch, O.random_scalars(0, n, 21)), _dev(torch, O.random_scalars(0, n, 22))
    out = torch.zeros(4, dtype=torch.int64, device="cuda")
    first = _call(hb, torch, 0, a, b, n, out)                        # set-up call: symbol look-up, kernels loaded
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert (_call(hb, torch, 0, a, b, n, out) == first).all()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "device memory in use changed across halo_scalar_dot_noalloc_dev calls"
    # the capacity itself is accepted by the size check (rejected only beyond it); pointers are not touched on rejection
    rc = lib.halo_scalar_dot_noalloc_dev(0, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr()),
                                         ctypes.c_size_t(MAX_N + 1), ctypes.c_void_p(out.data_ptr()))
    assert rc == HALO_EINVAL
    assert b"2^24" in lib.halo_last_error()
    with pytest.raises(hb.PreconditionError):
        _call(hb, torch, 2, a, b, n, out)                            # no such field
    # and the library still works afterwards
    assert (_call(hb, torch, 0, a, b, n, out) == first).all()


def test_full_capacity_uses_every_scratch_slot(hb):
    """n = 2^24: all 4096 slots.  Any 256-bit value below the modulus is a valid Montgomery representative, so random
    words with the top limb below 2^62 are uniform-enough field elements."""
    import torch
    rng = np.random.default_rng(2024)
    a = rng.integers(0, 1 << 63, size=(MAX_N, 4), dtype=np.uint64)
    b = rng.integers(0, 1 << 63, size=(MAX_N, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 62) - 1)
    b[:, 3] &= np.uint64((1 << 62) - 1)
    exp = hb.group.scalar_dot(1, a, b)
    out = torch.zeros(4, dtype=torch.int64, device="cuda")
    da, db = _dev(torch, a), _dev(torch, b)
    assert (_call(hb, torch, 1, da, db, MAX_N, out) == exp).all()
    # a short call right after the full one reads one slot only
    exp_s = hb.group.scalar_dot(1, a[:77], b[:77])
    assert (_call(hb, torch, 1, da, db, 77, out) == exp_s).all()
