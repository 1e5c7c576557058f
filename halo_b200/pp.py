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


CURVE_DIRS = ("pallas", "vesta")
G_BLOCKS_NO = 64            # crates/group/src/consts.rs:1-3: N = 2^20 generators in 64 blocks of 16 384


def decode_block(data, with_length=True, max_points=None):
    """One gs-NN.bin (with_length) or sh.bin (two points, no length) of the reference, decoded on the host:
    bincode 2 `standard()` varints -> (k, 8) Montgomery limb octets (x, y)."""
    data = bytes(data)
    cap = max_points if max_points is not None else (len(data) // 8 + 1 if with_length else 2)
    out = np.zeros((max(cap, 1), 8), dtype=np.uint64)
    got = ctypes.c_size_t(0)
    check(lib.halo_pp_decode_block(data, ctypes.c_size_t(len(data)), ctypes.c_int(1 if with_length else 0),
                                   ctypes.c_size_t(cap), ptr(out.reshape(-1)), ctypes.byref(got)))
    return out[:got.value].copy()


def new(curve, n, precompute_dir):
    """≙ PublicParams::new(n) (pp.rs:26-61) on the reference's own files: precompute_dir is crates/group/.precompute
    (or one curve's subdirectory of it).  The blocks are decoded inside the library and made resident."""
    import os
    _lib.ensure_init()
    d = precompute_dir
    if os.path.isdir(os.path.join(d, CURVE_DIRS[curve])):
        d = os.path.join(d, CURVE_DIRS[curve])
    need = min(G_BLOCKS_NO, (n + 16383) // 16384)
    blobs = [open(os.path.join(d, f"gs-{i:02d}.bin"), "rb").read() for i in range(need)]
    sh = open(os.path.join(d, "sh.bin"), "rb").read()
    return load_blobs(curve, blobs, sh, n)


def load_blobs(curve, blobs, sh, n):
    _lib.ensure_init()
    arr = (ctypes.c_char_p * len(blobs))(*blobs)
    lens = (ctypes.c_size_t * len(blobs))(*[len(b) for b in blobs])
    check(lib.halo_pp_load_blobs(curve, arr, lens, ctypes.c_size_t(len(blobs)), sh, ctypes.c_size_t(len(sh)), ctypes.c_size_t(n)))
    if curve not in _host:
        S, H = decode_block(sh, with_length=False)
        _host[curve] = PublicParams(curve, n, S.copy(), H.copy())
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
