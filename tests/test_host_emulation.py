"""The device limb algorithms (halo_b200/csrc/field.cuh, curve.cuh), compiled for the host with emulated PTX carry
flags, against the oracle.  Also the product's plain host path (used by the host-side transcript).  No GPU."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
_u64p = ctypes.POINTER(ctypes.c_uint64)


def _build(tag, flags):
    so = os.path.join(HERE, "emu", f"libemu_{tag}.so")
    src = os.path.join(HERE, "emu", "emu_lib.cpp")
    deps = [src] + [os.path.join(ROOT, "halo_b200/csrc", h) for h in ("field.cuh", "curve.cuh", "transcript.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I", os.path.join(ROOT, "halo_b200/csrc"),
                               *flags, "-o", so, src])
    return ctypes.CDLL(so)


# "plain" takes the BMI2 + ADX multiplication of the host path when the CPU has it, "plain_noadx" the generic one
@pytest.fixture(scope="module", params=["ptx", "plain", "plain_noadx"])
def emu(request):
    flags = {"ptx": [], "plain": ["-DEMU_PLAIN_HOST"], "plain_noadx": ["-DEMU_PLAIN_HOST", "-DHALO_NO_HOST_ADX"]}
    return _build(request.param, flags[request.param])


def _p(a):
    return a.ctypes.data_as(_u64p)


@pytest.mark.parametrize("field", [0, 1])
def test_field_ops(emu, oracle, field):
    O = oracle
    p = O.MODULUS[field]
    n = 3000
    a = O.random_scalars(field, n, 101)
    b = O.random_scalars(field, n, 202)
    edge = O.ints_to_mont(field, [0, 1, p - 1, p - 2, 2, (1 << 254) - 1, 1 << 32, (1 << 32) - 1])
    a[: len(edge)] = edge
    b[len(edge) // 2: len(edge) // 2 + len(edge)] = edge
    for op, ref in ((0, O.field_mul), (1, O.field_add), (2, O.field_sub)):
        out = np.empty_like(a)
        emu.emu_field_op(field, op, _p(a), _p(b), _p(out), ctypes.c_size_t(n))
        assert (out == ref(field, a, b)).all(), op
    out = np.empty_like(a)
    emu.emu_field_op(field, 6, _p(a), _p(b), _p(out), ctypes.c_size_t(n))          # dedicated squaring
    assert (out == O.field_mul(field, a, a)).all()
    emu.emu_field_op(field, 3, _p(a[:50].copy()), _p(b), _p(out), ctypes.c_size_t(50))
    assert (out[:50] == O.field_inv(field, a[:50])).all()
    emu.emu_field_op(field, 4, _p(a), _p(b), _p(out), ctypes.c_size_t(n))
    assert (out == O.field_to_canonical(field, a)).all()
    back = np.empty_like(a)
    emu.emu_field_op(field, 5, _p(out), _p(b), _p(back), ctypes.c_size_t(n))
    assert (back == a).all()


@pytest.mark.parametrize("field", [0, 1])
def test_host_sponge_and_lazy_dot(emu, oracle, field):
    """The product's host Poseidon sponge (transcript.cuh, with its one-reduction MDS rows) against the KAT-pinned
    oracle sponge, and the three-term dot product on edge values (all operands p-1: the largest accumulator)."""
    O = oracle
    p = O.MODULUS[field]
    for n in (0, 1, 2, 3, 5, 8):
        x = O.random_scalars(field, max(n, 1), 300 + n)[:n]
        out = np.empty(4, dtype=np.uint64)
        emu.emu_poseidon_hash(field, _p(np.ascontiguousarray(x)), ctypes.c_size_t(n), _p(out))
        assert (out == O.poseidon_hash(field, x)).all(), n
    n = 3 * 400
    a, b = O.random_scalars(field, n, 11), O.random_scalars(field, n, 12)
    edge = O.ints_to_mont(field, [p - 1] * 6 + [0, 1, p - 1, p - 1, 0, 0])
    a[: len(edge)] = edge
    b[: len(edge)] = O.ints_to_mont(field, [p - 1] * 6 + [p - 1, p - 1, 1, 0, 0, 0])
    out = np.empty((n // 3, 4), dtype=np.uint64)
    emu.emu_dot3(field, _p(a), _p(b), _p(out), ctypes.c_size_t(n))
    prod = O.field_mul(field, a, b).reshape(-1, 3, 4)
    exp = O.field_add(field, O.field_add(field, prod[:, 0], prod[:, 1]), prod[:, 2])
    assert (out == exp).all()


@pytest.mark.parametrize("curve", [0, 1])
def test_curve_ops(emu, oracle, curve):
    O = oracle
    pts = O.points_arith(curve, 40)
    # bucket-style accumulation incl. duplicates (doubling), inverse pairs (-> identity) and identity inputs
    seq = [pts[0], pts[1], pts[1], pts[2], pts[0], np.zeros(8, dtype=np.uint64), pts[5], pts[5], pts[5]]
    neg = [0, 0, 0, 1, 1, 0, 0, 1, 0]
    arr = np.ascontiguousarray(np.stack(seq))
    out = np.empty(12, dtype=np.uint64)
    emu.emu_madd_chain(curve, _p(arr), (ctypes.c_int * len(neg))(*neg), ctypes.c_size_t(len(seq)), 2, _p(out))
    sfield = O.SCALAR_FIELD[curve]
    # expected: 4 * (2*P1 - P2 + P5)
    exp = O.msm_naive(curve, np.stack([pts[1], pts[2], pts[5]]), O.ints_to_mont(sfield, [8, O.MODULUS[sfield] - 4, 4]))
    assert (O.point_to_affine(curve, out) == O.point_to_affine(curve, exp)).all()
    # P + (-P) straight away, then continue
    seq2 = np.ascontiguousarray(np.stack([pts[3], pts[3], pts[7]]))
    emu.emu_madd_chain(curve, _p(seq2), (ctypes.c_int * 3)(0, 1, 0), ctypes.c_size_t(3), 0, _p(out))
    assert (O.point_to_affine(curve, out)[0] == pts[7]).all()
    # general add, to-affine, scalar mul
    a = O.point_mul(curve, O.generator(curve), O.ints_to_mont(sfield, [123456789])[0])
    b = O.point_mul(curve, O.generator(curve), O.ints_to_mont(sfield, [987654321])[0])
    o_xyz, o_xy = np.empty(12, dtype=np.uint64), np.empty(8, dtype=np.uint64)
    for x, y in ((a, b), (a, a), (a, O.affine_to_jac(curve, np.zeros(8, dtype=np.uint64))[0])):
        emu.emu_add(curve, _p(np.ascontiguousarray(x)), _p(np.ascontiguousarray(y)), _p(o_xyz), _p(o_xy))
        exp = O.point_to_affine(curve, O.point_add(curve, x, y))[0]
        assert (o_xy == exp).all() and (O.point_to_affine(curve, o_xyz)[0] == exp).all()
    k = O.random_scalars(sfield, 1, 77)
    kc = O.field_to_canonical(sfield, k)
    emu.emu_mul(curve, _p(np.ascontiguousarray(a)), _p(kc), _p(o_xyz))
    assert (O.point_to_affine(curve, o_xyz) == O.point_to_affine(curve, O.point_mul(curve, a, k[0]))).all()


@pytest.mark.parametrize("field", [0, 1])
def test_safegcd_inversion(emu, oracle, field):
    """fe_inv_safegcd (divsteps on 30-bit limbs, the inversion of the batched-affine bucket accumulation) against the
    oracle's Fermat inverse: random elements, the edges of the range, small values and powers of two; 0 -> 0."""
    O = oracle
    p = O.MODULUS[field]
    n = 4000
    a = O.random_scalars(field, n, 777 + field)
    ints = [0, 1, 2, 3, p - 1, p - 2, (p - 1) // 2, (p + 1) // 2, (1 << 254) - 1, 1 << 253, (1 << 30) - 1, 1 << 30, (1 << 30) + 1]
    ints += [1 << k for k in range(0, 254, 7)] + [p - (1 << k) for k in range(0, 254, 11)]
    edge = O.ints_to_mont(field, ints)
    raw = np.array([O.int_to_limbs(x) for x in ints], dtype=np.uint64)      # the same integers as raw limb patterns
    a[: len(edge)] = edge
    a[len(edge): 2 * len(edge)] = raw
    out = np.empty_like(a)
    emu.emu_field_op(field, 7, _p(a), _p(a), _p(out), ctypes.c_size_t(n))
    assert (out == O.field_inv(field, a)).all(), np.flatnonzero((out != O.field_inv(field, a)).any(axis=1))[:10]
    one = O.ints_to_mont(field, [1])[0]
    nz = a.any(axis=1)
    assert (O.field_mul(field, a[nz], out[nz]) == one).all()
    assert (out[~nz] == 0).all()
