"""crates/accumulation/src/acc.rs — AS_DL: prover / verifier / decider over PC_DL instances."""
import ctypes

import numpy as np

from . import _lib, pcdl
from ._lib import lib, check, ptr


class Accumulator:
    """≙ acc::Accumulator { q: Instance } (acc.rs:24-58)."""

    def __init__(self, q):
        self.q = q

    @classmethod
    def zero(cls, curve, n):
        """≙ Accumulator::zero(n, 1) (acc.rs:37-41): accumulate one zero instance."""
        return prover(curve, [pcdl.Instance.zero(curve, n)])


def _pack(qs):
    assert len({q.w.shape[0] for q in qs}) <= 1, "instances of one accumulation share d"
    return np.ascontiguousarray(np.stack([q.w for q in qs])) if qs else np.zeros((0, 0), dtype=np.uint64)


def prover(curve, qs):
    """≙ acc::prover (acc.rs:178-198); raises Rejected when an input instance fails its succinct check."""
    _lib.ensure_init()
    flat = _pack(qs)
    out = np.zeros(flat.shape[1] if qs else 0, dtype=np.uint64)
    check(lib.halo_acc_prover(curve, ptr(flat.reshape(-1)) if qs else None, ctypes.c_size_t(len(qs)), ptr(out)))
    return Accumulator(pcdl.Instance(curve, out))


def verifier(curve, qs, acc):
    """≙ acc::verifier (acc.rs:200-213)."""
    _lib.ensure_init()
    flat = _pack(qs)
    check(lib.halo_acc_verifier(curve, ptr(flat.reshape(-1)) if qs else None, ctypes.c_size_t(len(qs)), ptr(acc.q.w)))


def decider(curve, acc):
    """≙ acc::decider (acc.rs:215-217) = pcdl::check of the accumulator."""
    _lib.ensure_init()
    check(lib.halo_acc_decider(curve, ptr(acc.q.w)))
