"""ark-serialize compressed wire formats of Instance / EvalProof / Accumulator (crates/accumulation/src/pcdl.rs:26,159,
acc.rs:22) and of the benchmark file qs.bin = Vec<(usize, Instance, Accumulator)> (crates/accumulation/src/main.rs:59-91),
over the C ABI (halo_instance_serialize / halo_instance_deserialize ...).  Host only."""
import ctypes
import struct

import numpy as np

from . import pcdl
from . import acc as _acc
from ._lib import lib, check, ptr, as_u64

_u8p = ctypes.POINTER(ctypes.c_uint8)


def point_compress(curve, xyz):
    out = (ctypes.c_uint8 * 33)()
    check(lib.halo_point_compress(curve, ptr(as_u64(xyz, 12).reshape(12)), out))
    return bytes(out)


def point_decompress(curve, data):
    buf = (ctypes.c_uint8 * 33).from_buffer_copy(bytes(data[:33]))
    out = np.empty(12, dtype=np.uint64)
    check(lib.halo_point_decompress(curve, buf, ptr(out)))
    return out


def scalar_serialize(field, s):
    out = (ctypes.c_uint8 * 32)()
    check(lib.halo_scalar_serialize(field, ptr(as_u64(s, 4).reshape(4)), out))
    return bytes(out)


def scalar_deserialize(field, data):
    buf = (ctypes.c_uint8 * 32).from_buffer_copy(bytes(data[:32]))
    out = np.empty(4, dtype=np.uint64)
    check(lib.halo_scalar_deserialize(field, buf, ptr(out)))
    return out


def instance_serialize(q):
    """Instance::serialize_compressed -> bytes."""
    need = ctypes.c_size_t(0)
    lib.halo_instance_serialize(q.curve, ptr(q.w), None, ctypes.c_size_t(0), ctypes.byref(need))
    out = (ctypes.c_uint8 * need.value)()
    check(lib.halo_instance_serialize(q.curve, ptr(q.w), out, ctypes.c_size_t(need.value), ctypes.byref(need)))
    return bytes(out)


def instance_deserialize(curve, data, offset=0):
    """Instance::deserialize_compressed from data[offset:] -> (Instance, bytes consumed)."""
    data = bytes(data)
    if len(data) - offset < 41:
        raise pcdl._lib.PreconditionError(pcdl._lib.HALO_EINVAL, "truncated")
    d = struct.unpack_from("<Q", data, offset + 33)[0]
    lg = max((d + 1).bit_length() - 1, 0) if d + 1 < (1 << 41) else 0
    words = np.zeros(24 * lg + 54, dtype=np.uint64)
    buf = (ctypes.c_uint8 * (len(data) - offset)).from_buffer_copy(data[offset:])
    used = ctypes.c_size_t(0)
    check(lib.halo_instance_deserialize(curve, buf, ctypes.c_size_t(len(data) - offset), ptr(words), ctypes.c_size_t(words.shape[0]),
                                        ctypes.byref(used)))
    return pcdl.Instance(curve, words), used.value


def accumulator_serialize(a):
    return instance_serialize(a.q)


def accumulator_deserialize(curve, data, offset=0):
    q, used = instance_deserialize(curve, data, offset)
    return _acc.Accumulator(q), used


def qs_serialize(entries):
    """Vec<(usize, Instance, Accumulator)>::serialize_compressed (accumulation/src/main.rs:84-86)."""
    out = [struct.pack("<Q", len(entries))]
    for n, q, a in entries:
        out += [struct.pack("<Q", n), instance_serialize(q), accumulator_serialize(a)]
    return b"".join(out)


def qs_deserialize(curve, data):
    data = bytes(data)
    (cnt,) = struct.unpack_from("<Q", data, 0)
    pos, out = 8, []
    for _ in range(cnt):
        (n,) = struct.unpack_from("<Q", data, pos)
        pos += 8
        q, used = instance_deserialize(curve, data, pos)
        pos += used
        a, used = accumulator_deserialize(curve, data, pos)
        pos += used
        out.append((n, q, a))
    if pos != len(data):
        raise pcdl._lib.PreconditionError(pcdl._lib.HALO_EINVAL, "trailing bytes")
    return out
