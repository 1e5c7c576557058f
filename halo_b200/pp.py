"""PublicParams (crates/group/src/pp.rs:10-94): per-curve generator set, resident in HBM once per process."""
import ctypes

import numpy as np

from . import _lib
from ._lib import lib, check, ptr, as_u64

_host = {}


class PublicParams:
    """Host view of what was uploaded (Gs stays on the device; S and H are kept for the host-side transcript)."""

    def __init__(self, curve, n, S, H):
        self.curve, self.n, self.S, self.H = curve, n, S, H

    def __len__(self):
        return self.n

    @property
    def D(self):
        return self.n - 1


def set_pp(curve, gs_xy, S_xy, H_xy):
    """≙ PublicParams::set_pp (pp.rs:63-82): idempotent, monotone; the caller supplies the decoded generators
    (the reference embeds them as bincode blobs, pp.rs:36-53)."""
    _lib.ensure_init()
    gs = as_u64(gs_xy, 8)
    S = as_u64(S_xy, 8).reshape(8)
    H = as_u64(H_xy, 8).reshape(8)
    check(lib.halo_pp_load(curve, ptr(gs), ctypes.c_size_t(gs.shape[0]), ptr(S), ptr(H)))
    if curve not in _host:
        _host[curve] = PublicParams(curve, gs.shape[0], S.copy(), H.copy())
    return _host[curve]


def get_pp(curve):
    """≙ PublicParams::get_pp (pp.rs:84-94).  The blobs live with the Rust crate, so unlike the reference this
    cannot lazily load 2^20 points: it raises if set_pp was never called."""
    if curve not in _host or lib.halo_pp_len(curve) == 0:
        raise _lib.ParamsError(_lib.HALO_EPARAMS, "public parameters not loaded for this curve")
    return _host[curve]


def pp_len(curve):
    return int(lib.halo_pp_len(curve))


def reset(curve):
    _lib.ensure_init()
    check(lib.halo_pp_reset(curve))
    _host.pop(curve, None)
