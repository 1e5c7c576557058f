"""Poseidon on the device (include/halo_b200.h, halo_poseidon_hash_batch): many independent sponges at once.
≙ crates/poseidon/src/inner_sponge.rs:11-115 (InnerSponge::new / absorb / squeeze, Kimchi parameters) for batches; the
single serial Fiat-Shamir transcript of a proof stays on the host inside the library."""
import ctypes

import numpy as np

from . import _lib
from ._lib import lib, check, ptr


def hash_batch(field, inputs):
    """inputs: (count, len, 4) Montgomery elements -> (count, 4): sponge k absorbs inputs[k] and squeezes one element."""
    _lib.ensure_init()
    a = np.ascontiguousarray(inputs, dtype=np.uint64)
    assert a.ndim == 3 and a.shape[2] == 4
    count, length = a.shape[0], a.shape[1]
    out = np.zeros((count, 4), dtype=np.uint64)
    check(lib.halo_poseidon_hash_batch(field, ptr(a.reshape(-1)) if a.size else None, ctypes.c_size_t(length),
                                       ctypes.c_size_t(count), ptr(out.reshape(-1)) if count else None))
    return out
