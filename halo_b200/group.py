"""crates/group/src/group.rs:43-66 — point_dot_affine / point_dot on the GPU."""
import ctypes

import numpy as np

from . import _lib
from ._lib import lib, check, ptr, as_u64


def point_dot_affine(curve, xs, Gs):
    """≙ point_dot_affine (group.rs:48-50): MSM over min(len) pairs, no length check ("unchecked")."""
    _lib.ensure_init()
    xs, Gs = as_u64(xs, 4), as_u64(Gs, 8)
    k = min(xs.shape[0], Gs.shape[0])
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_msm(curve, ptr(Gs), ctypes.c_size_t(0), ptr(xs), ctypes.c_size_t(k), ptr(out)))
    return out


def point_dot(curve, xs, Gs_proj):
    """≙ point_dot (group.rs:53-56): projective bases."""
    _lib.ensure_init()
    xs, Gs = as_u64(xs, 4), as_u64(Gs_proj, 12)
    k = min(xs.shape[0], Gs.shape[0])
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_msm_proj(curve, ptr(Gs), ptr(xs), ctypes.c_size_t(k), ptr(out)))
    return out


def msm_resident(curve, xs, base_off=0):
    """MSM against the resident generators Gs[base_off : base_off+len(xs)]."""
    _lib.ensure_init()
    xs = as_u64(xs, 4)
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_msm(curve, None, ctypes.c_size_t(base_off), ptr(xs), ctypes.c_size_t(xs.shape[0]), ptr(out)))
    return out


def field_op(field, op, a, b=None):
    """Elementwise field arithmetic on the GPU (test hook). op: mul add sub inv to_canonical from_canonical."""
    _lib.ensure_init()
    code = {"mul": 0, "add": 1, "sub": 2, "inv": 3, "to_canonical": 4, "from_canonical": 5}[op]
    a = as_u64(a, 4)
    b = as_u64(b, 4) if b is not None else None
    out = np.empty_like(a)
    check(lib.halo_field_op(field, code, ptr(a), ptr(b), ptr(out), ctypes.c_size_t(a.shape[0])))
    return out


def msm_batch_resident(curve, xs, base_off=0):
    """b MSMs over the same resident generators Gs[base_off : base_off+k] in one device pipeline — the shape of the
    16 commits of a prover round (crates/plonk/src/plonk/protocol.rs:114,263).  xs: (b, k, 4)."""
    _lib.ensure_init()
    xs = np.ascontiguousarray(xs, dtype=np.uint64)
    assert xs.ndim == 3 and xs.shape[2] == 4
    b, k = xs.shape[0], xs.shape[1]
    out = np.empty((b, 12), dtype=np.uint64)
    check(lib.halo_msm_batch(curve, ctypes.c_size_t(base_off), ptr(xs.reshape(-1, 4)), ctypes.c_size_t(k),
                             ctypes.c_size_t(b), ptr(out)))
    return out


def scalar_dot(field, xs, ys):
    """≙ scalar_dot (group.rs:43-45): sum_i xs[i]*ys[i] over min(len) pairs (zip semantics)."""
    _lib.ensure_init()
    xs, ys = as_u64(xs, 4), as_u64(ys, 4)
    n = min(xs.shape[0], ys.shape[0])
    out = np.empty(4, dtype=np.uint64)
    check(lib.halo_scalar_dot(field, ptr(xs) if n else None, ptr(ys) if n else None, ctypes.c_size_t(n), ptr(out)))
    return out


def construct_powers(field, z, n):
    """≙ construct_powers (group.rs:58-66): (1, z, ..., z^(n-1))."""
    _lib.ensure_init()
    out = np.empty((n, 4), dtype=np.uint64)
    check(lib.halo_construct_powers(field, ptr(as_u64(z, 4).reshape(4)), ctypes.c_size_t(n), ptr(out) if n else None))
    return out
