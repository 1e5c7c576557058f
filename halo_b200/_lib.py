"""ctypes binding of libhalo_b200.so (the C ABI in include/halo_b200.h).

There is no CPU fallback: if the CUDA library is missing or cannot be loaded the import fails loudly.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libhalo_b200.so")

HALO_OK, HALO_EINVAL, HALO_ECUDA, HALO_ENOINIT, HALO_EPARAMS, HALO_EREJECT = 0, -1, -2, -3, -4, -5


class HaloError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"halo_b200 error {code}: {msg}")
        self.code = code


class PreconditionError(HaloError):
    """The reference would assert!/panic! here."""


class ParamsError(HaloError):
    """Public parameters missing / too small (crates/group/src/pp.rs:27-32,63-82)."""


class Rejected(HaloError):
    """A verifier rejected (anyhow::ensure! in the reference)."""


if not os.path.exists(SO_PATH):
    raise ImportError(
        f"{SO_PATH} is missing: build it with `make -C halo_b200/csrc` (or __graft_entry__.build()). "
        "halo_b200 has no CPU fallback.")
lib = ctypes.CDLL(SO_PATH)

u64p = ctypes.POINTER(ctypes.c_uint64)
lib.halo_last_error.restype = ctypes.c_char_p
lib.halo_pp_len.restype = ctypes.c_size_t


def check(rc):
    if rc == HALO_OK:
        return
    msg = lib.halo_last_error().decode(errors="replace")
    cls = {HALO_EINVAL: PreconditionError, HALO_EPARAMS: ParamsError, HALO_EREJECT: Rejected}.get(rc, HaloError)
    raise cls(rc, msg)


def ptr(a):
    if a is None:
        return None
    assert isinstance(a, np.ndarray) and a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"], "need contiguous uint64 array"
    return a.ctypes.data_as(u64p)


def as_u64(a, last):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    assert a.size % last == 0
    return a.reshape(-1, last)


_inited = False


def init(device=-1):
    global _inited
    check(lib.halo_init(ctypes.c_int(device)))
    _inited = True


def ensure_init():
    if not _inited:
        init(-1)
