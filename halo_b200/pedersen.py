"""crates/accumulation/src/pedersen.rs:7-27 — Pedersen vector commitment."""
import numpy as np

from . import _lib, consts, group, pp as _pp
from ._lib import as_u64


def commit(curve, w, Gs, ms):
    """≙ pedersen::commit(w, Gs, ms) = MSM(Gs, ms) [+ w*S].

    `Gs` is either an (k,8) array of affine points or a (offset, length) slice of the resident generators
    (the reference always passes `&pp.Gs[0..n]`).  Panics -> PreconditionError when len(Gs) < len(ms)
    (pedersen.rs:14-19)."""
    ms = as_u64(ms, 4)
    if isinstance(Gs, tuple):
        off, length = Gs
        if length < ms.shape[0]:
            raise _lib.PreconditionError(_lib.HALO_EINVAL, f"ms must be larger than Gs: (Gs: {length}), (ms: {ms.shape[0]})")
        acc = group.msm_resident(curve, ms, off)
    else:
        Gs = as_u64(Gs, 8)
        if Gs.shape[0] < ms.shape[0]:
            raise _lib.PreconditionError(_lib.HALO_EINVAL, f"ms must be larger than Gs: (Gs: {Gs.shape[0]}), (ms: {ms.shape[0]})")
        acc = group.point_dot_affine(curve, ms, Gs)
    if w is None:
        return acc
    # pp.S * w + acc (pedersen.rs:22-23) as a 2-term projective MSM so the arithmetic stays on the device
    P = _pp.get_pp(curve)
    bases = np.concatenate([consts.affine_to_projective(curve, P.S), acc.reshape(1, 12)])
    scalars = np.stack([as_u64(w, 4).reshape(4), consts.one(consts.SCALAR_FIELD[curve])])
    return group.point_dot(curve, scalars, bases)
