"""halo_b200 — B200-native MSM / IPA / NTT hot path for rasmus-kirk/halo.

Python host-side mirror of the reference's Rust interface for this path (crates/group, crates/accumulation),
over the C ABI of libhalo_b200.so.  Elements are numpy uint64 arrays holding the reference's in-memory form
(Montgomery limbs): scalars (...,4), affine points (...,8) with identity (0,0), projective points (...,12).
"""
from . import _lib
from ._lib import HaloError, ParamsError, PreconditionError, Rejected, init
from . import consts, group, pp, pedersen, poly, pcdl, acc, plonk, dist, serialize, poseidon

PALLAS, VESTA = 0, 1
FP, FQ = 0, 1

__all__ = ["init", "group", "pp", "pedersen", "poly", "pcdl", "acc", "plonk", "dist", "serialize", "poseidon", "consts", "HaloError", "ParamsError", "PreconditionError", "Rejected",
           "PALLAS", "VESTA", "FP", "FQ"]
