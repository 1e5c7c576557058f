"""ctypes binding of the CPU oracle (oracle/liboracle.so) — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.  The product package (halo_b200/) never does.
All elements are numpy uint64 arrays of Montgomery limbs: scalars (...,4),
affine points (...,8) with identity = all-zero, Jacobian points (...,12).
"""
import ctypes
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "liboracle.so")

PALLAS, VESTA = 0, 1
FP, FQ = 0, 1          # halo naming: Fp = Pallas scalar field, Fq = Pallas base field
SCALAR_FIELD = {PALLAS: FP, VESTA: FQ}
BASE_FIELD = {PALLAS: FQ, VESTA: FP}

MODULUS = {
    FP: 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001,
    FQ: 0x40000000000000000000000000000000224698fc094cf91b992d30ed00000001,
}
R = 1 << 256


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(_ROOT, "oracle")])


def _load():
    if not os.path.exists(_SO):
        build()
    return ctypes.CDLL(_SO)


_lib = _load()
_u64p = ctypes.POINTER(ctypes.c_uint64)
_lib.orc_last_error.restype = ctypes.c_char_p
_lib.orc_bincode_decode_u64s.restype = ctypes.c_long
_lib.orc_instance_words.restype = ctypes.c_size_t


def _p(a):
    if a is None:
        return None
    assert a.dtype == np.uint64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_u64p)


def _c(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def set_threads(n):
    _lib.orc_set_threads(ctypes.c_uint(n))


def last_error():
    return _lib.orc_last_error().decode()


# ---------------- integer <-> limb helpers (pure python, independent of the .so) -------------
def int_to_limbs(x):
    return [(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)]


def limbs_to_int(l):
    return sum(int(v) << (64 * i) for i, v in enumerate(l))


def to_mont(field, x):
    return np.array(int_to_limbs((x * R) % MODULUS[field]), dtype=np.uint64)


def from_mont(field, limbs):
    return (limbs_to_int(limbs) * pow(R, -1, MODULUS[field])) % MODULUS[field]


def ints_to_mont(field, xs):
    return np.array([int_to_limbs((x * R) % MODULUS[field]) for x in xs], dtype=np.uint64).reshape(-1, 4)


# ---------------- field ----------------
def _binop(fn, field, a, b):
    a, b = _c(a).reshape(-1, 4), _c(b).reshape(-1, 4)
    out = np.empty_like(a)
    fn(field, _p(a), _p(b), _p(out), ctypes.c_size_t(a.shape[0]))
    return out


def field_mul(field, a, b):
    return _binop(_lib.orc_field_mul, field, a, b)


def field_add(field, a, b):
    return _binop(_lib.orc_field_add, field, a, b)


def field_sub(field, a, b):
    return _binop(_lib.orc_field_sub, field, a, b)


def field_inv(field, a):
    a = _c(a).reshape(-1, 4)
    out = np.empty_like(a)
    _lib.orc_field_inv(field, _p(a), _p(out), ctypes.c_size_t(a.shape[0]))
    return out


def field_to_canonical(field, a):
    a = _c(a).reshape(-1, 4)
    out = np.empty_like(a)
    _lib.orc_field_to_canonical(field, _p(a), _p(out), ctypes.c_size_t(a.shape[0]))
    return out


def field_from_canonical(field, a):
    a = _c(a).reshape(-1, 4)
    out = np.empty_like(a)
    _lib.orc_field_from_canonical(field, _p(a), _p(out), ctypes.c_size_t(a.shape[0]))
    return out


def root_of_unity(field, log_n):
    out = np.empty(4, dtype=np.uint64)
    _lib.orc_root_of_unity(field, ctypes.c_uint(log_n), _p(out))
    return out


# ---------------- deterministic inputs ----------------
def _splitmix64(state):
    state = (state + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
    z = state
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    return state, z ^ (z >> 31)


def random_scalars(field, n, seed):
    """Uniform elements of `field` in Montgomery form.  Vectorised SplitMix64 stream,
    top limb masked to 62 bits then rejection against the modulus is unnecessary for the
    low 254 bits except for the tiny band [2^254, p): we draw 254-bit values, which are
    all < p, and feed them in AS Montgomery representatives (a bijection of [0,p) minus a
    2^-128 fraction), so the canonical scalars are uniform up to that negligible band."""
    idx = np.arange(n * 4, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = (idx + np.uint64(1)) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(seed)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    out = z.reshape(n, 4).copy()
    out[:, 3] &= np.uint64((1 << 62) - 1)
    return out


def edge_scalars(field, n, seed):
    """Stress distribution: 1/16 each of {0, 1, p-1, < 2^16, duplicate of previous}."""
    out = random_scalars(field, n, seed)
    p = MODULUS[field]
    one = to_mont(field, 1)
    pm1 = to_mont(field, p - 1)
    for i in range(n):
        k = (i * 2654435761 + seed) % 16
        if k == 0:
            out[i] = 0
        elif k == 1:
            out[i] = one
        elif k == 2:
            out[i] = pm1
        elif k == 3:
            out[i] = to_mont(field, (i * 7919 + 1) % 65536)
        elif k == 4 and i > 0:
            out[i] = out[i - 1]
    return out


# ---------------- curve ----------------
def point_to_affine(curve, xyz):
    xyz = _c(xyz).reshape(-1, 12)
    out = np.empty((xyz.shape[0], 8), dtype=np.uint64)
    _lib.orc_point_to_affine(curve, _p(xyz), _p(out), ctypes.c_size_t(xyz.shape[0]))
    return out


def affine_to_jac(curve, xy):
    xy = _c(xy).reshape(-1, 8)
    out = np.zeros((xy.shape[0], 12), dtype=np.uint64)
    one = to_mont(BASE_FIELD[curve], 1)
    for i in range(xy.shape[0]):
        if not xy[i].any():
            out[i, 0:4] = one
            out[i, 4:8] = one
        else:
            out[i, 0:8] = xy[i]
            out[i, 8:12] = one
    return out


def point_on_curve(curve, xy):
    return bool(_lib.orc_point_on_curve(curve, _p(_c(xy).reshape(8))))


def point_add(curve, a, b):
    out = np.empty(12, dtype=np.uint64)
    _lib.orc_point_add(curve, _p(_c(a).reshape(12)), _p(_c(b).reshape(12)), _p(out))
    return out


def point_mul(curve, a, s):
    out = np.empty(12, dtype=np.uint64)
    _lib.orc_point_mul(curve, _p(_c(a).reshape(12)), _p(_c(s).reshape(4)), _p(out))
    return out


def generator(curve):
    out = np.empty(12, dtype=np.uint64)
    _lib.orc_generator(curve, _p(out))
    return out


def points_arith(curve, n, k0=0x1234567, kd=0x9E3779B9):
    """n distinct affine points P_i = k0*G + i*kd*G."""
    out = np.empty((n, 8), dtype=np.uint64)
    _lib.orc_points_arith(curve, ctypes.c_uint64(k0), ctypes.c_uint64(kd), ctypes.c_size_t(n), _p(out))
    return out


# ---------------- MSM / NTT / Poseidon ----------------
def msm(curve, bases_xy, scalars):
    bases_xy, scalars = _c(bases_xy).reshape(-1, 8), _c(scalars).reshape(-1, 4)
    k = min(bases_xy.shape[0], scalars.shape[0])
    out = np.empty(12, dtype=np.uint64)
    _lib.orc_msm(curve, _p(bases_xy), _p(scalars), ctypes.c_size_t(k), _p(out))
    return out


def msm_naive(curve, bases_xy, scalars):
    bases_xy, scalars = _c(bases_xy).reshape(-1, 8), _c(scalars).reshape(-1, 4)
    k = min(bases_xy.shape[0], scalars.shape[0])
    out = np.empty(12, dtype=np.uint64)
    _lib.orc_msm_naive(curve, _p(bases_xy), _p(scalars), ctypes.c_size_t(k), _p(out))
    return out


def msm_pp(curve, off, scalars):
    scalars = _c(scalars).reshape(-1, 4)
    out = np.empty(12, dtype=np.uint64)
    rc = _lib.orc_msm_pp(curve, ctypes.c_size_t(off), _p(scalars), ctypes.c_size_t(scalars.shape[0]), _p(out))
    assert rc == 0
    return out


def ntt(field, data, inverse=False, coset=None):
    data = _c(data).copy()
    shape = data.shape
    d2 = data.reshape(-1, shape[-2], 4) if data.ndim >= 2 else None
    n = d2.shape[1]
    log_n = n.bit_length() - 1
    assert 1 << log_n == n
    cs = _c(coset).reshape(4) if coset is not None else None
    _lib.orc_ntt(field, _p(d2), ctypes.c_uint(log_n), ctypes.c_size_t(d2.shape[0]), int(inverse), _p(cs))
    return d2.reshape(shape)


def dft_naive(field, data):
    data = _c(data).reshape(-1, 4)
    n = data.shape[0]
    out = np.empty_like(data)
    _lib.orc_dft_naive(field, _p(data), _p(out), ctypes.c_uint(n.bit_length() - 1))
    return out


def poseidon_hash(field, inputs):
    inputs = _c(inputs).reshape(-1, 4)
    out = np.empty(4, dtype=np.uint64)
    _lib.orc_poseidon_hash(field, _p(inputs), ctypes.c_size_t(inputs.shape[0]), _p(out))
    return out


# ---------------- public parameters ----------------
def pp_generate(curve, n):
    gs = np.empty((n, 8), dtype=np.uint64)
    s = np.empty(8, dtype=np.uint64)
    h = np.empty(8, dtype=np.uint64)
    _lib.orc_pp_generate(curve, ctypes.c_size_t(n), _p(gs), _p(s), _p(h))
    return gs, s, h


def pp_set(curve, gs, s, h):
    gs = _c(gs).reshape(-1, 8)
    _lib.orc_pp_set(curve, _p(gs), ctypes.c_size_t(gs.shape[0]), _p(_c(s).reshape(8)), _p(_c(h).reshape(8)))


def bincode_decode(blob, max_words, has_len_prefix=True):
    buf = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
    out = np.empty(max_words, dtype=np.uint64)
    n = _lib.orc_bincode_decode_u64s(buf, ctypes.c_size_t(len(blob)), _p(out), ctypes.c_size_t(max_words), int(has_len_prefix))
    assert n >= 0, "bincode decode failed"
    return out[:n]


# ---------------- PCDL / ASDL (flat instance layout, see include/halo_b200.h) ----------------
def instance_words(lg_n):
    return int(_lib.orc_instance_words(ctypes.c_uint(lg_n)))


class OracleError(Exception):
    pass


def pcdl_commit(curve, coeffs, d, w=None):
    coeffs = _c(coeffs).reshape(-1, 4)
    out = np.empty(12, dtype=np.uint64)
    rc = _lib.orc_pcdl_commit(curve, _p(coeffs), ctypes.c_size_t(coeffs.shape[0]), ctypes.c_size_t(d),
                              _p(_c(w).reshape(4)) if w is not None else None, _p(out))
    if rc != 0:
        raise OracleError(last_error())
    return out


def instance_open(curve, coeffs, d, z, w=None, hide_q=None, hide_wbar=None):
    coeffs = _c(coeffs).reshape(-1, 4)
    lg_n = (d + 1).bit_length() - 1
    out = np.zeros(instance_words(lg_n), dtype=np.uint64)
    rc = _lib.orc_instance_open(curve, _p(coeffs), ctypes.c_size_t(coeffs.shape[0]), ctypes.c_size_t(d),
                                _p(_c(z).reshape(4)),
                                _p(_c(w).reshape(4)) if w is not None else None,
                                _p(_c(hide_q).reshape(-1, 4)) if hide_q is not None else None,
                                _p(_c(hide_wbar).reshape(4)) if hide_wbar is not None else None,
                                _p(out))
    if rc != 0:
        raise OracleError(last_error())
    return out


def pcdl_succinct_check(curve, inst):
    """Returns (accept, xis, U)."""
    inst = _c(inst)
    d = int(inst[12])
    lg_n = (d + 1).bit_length() - 1
    xis = np.zeros((lg_n + 1, 4), dtype=np.uint64)
    U = np.zeros(12, dtype=np.uint64)
    rc = _lib.orc_pcdl_succinct_check(curve, _p(inst), ctypes.c_uint(lg_n), _p(xis), _p(U))
    if rc < 0:
        raise OracleError(last_error())
    return rc == 0, xis, U


def pcdl_check(curve, inst):
    rc = _lib.orc_pcdl_check(curve, _p(_c(inst)))
    if rc < 0:
        raise OracleError(last_error())
    return rc == 0


def hpoly_coeffs(scalar_field, xis):
    xis = _c(xis).reshape(-1, 4)
    lg_n = xis.shape[0] - 1
    out = np.empty((1 << lg_n, 4), dtype=np.uint64)
    _lib.orc_hpoly_coeffs(scalar_field, _p(xis), ctypes.c_uint(lg_n), _p(out))
    return out


def hpoly_eval(scalar_field, xis, z):
    xis = _c(xis).reshape(-1, 4)
    out = np.empty(4, dtype=np.uint64)
    _lib.orc_hpoly_eval(scalar_field, _p(xis), ctypes.c_uint(xis.shape[0] - 1), _p(_c(z).reshape(4)), _p(out))
    return out


def acc_prover(curve, insts, lg_n):
    insts = _c(insts).reshape(-1, instance_words(lg_n))
    out = np.zeros(instance_words(lg_n), dtype=np.uint64)
    rc = _lib.orc_acc_prover(curve, _p(insts), ctypes.c_size_t(insts.shape[0]), ctypes.c_uint(lg_n), _p(out))
    if rc != 0:
        raise OracleError(last_error())
    return out


def acc_verifier(curve, insts, lg_n, acc):
    insts = _c(insts).reshape(-1, instance_words(lg_n))
    rc = _lib.orc_acc_verifier(curve, _p(insts), ctypes.c_size_t(insts.shape[0]), ctypes.c_uint(lg_n), _p(_c(acc)))
    if rc < 0:
        raise OracleError(last_error())
    return rc == 0


def acc_decider(curve, acc):
    rc = _lib.orc_acc_decider(curve, _p(_c(acc)))
    if rc < 0:
        raise OracleError(last_error())
    return rc == 0


def fold_generators(curve, gl, gr, xi):
    gl, gr = _c(gl).reshape(-1, 8), _c(gr).reshape(-1, 8)
    out = np.empty_like(gl)
    _lib.orc_fold_generators(curve, _p(gl), _p(gr), _p(_c(xi).reshape(4)), ctypes.c_size_t(gl.shape[0]), _p(out))
    return out


# flat instance accessors
def inst_fields(inst):
    inst = np.asarray(inst, dtype=np.uint64)
    d = int(inst[12])
    L = (d + 1).bit_length() - 1
    o = 21
    f = {"C": inst[0:12], "d": d, "z": inst[13:17], "v": inst[17:21]}
    f["Ls"] = inst[o:o + 12 * L].reshape(L, 12); o += 12 * L
    f["Rs"] = inst[o:o + 12 * L].reshape(L, 12); o += 12 * L
    f["U"] = inst[o:o + 12]; o += 12
    f["c"] = inst[o:o + 4]; o += 4
    f["hiding"] = int(inst[o]); o += 1
    f["C_bar"] = inst[o:o + 12]; o += 12
    f["w_prime"] = inst[o:o + 4]
    return f


# ---------------- PLONK prover / verifier (oracle/plonk.hpp) ----------------
def plonk_proof_words(lg_n):
    return int(_lib.orc_plonk_proof_words(ctypes.c_uint(lg_n)))


def plonk_prove(curve, lg_n, polys, w_evals, public_inputs, acc_prev):
    """polys: (57, n, 4) coefficient columns in the order qs, ws, rs, ids, sigmas; w_evals: (3, n, 4)."""
    n = 1 << lg_n
    polys = _c(polys).reshape(57, n, 4)
    w_evals = _c(w_evals).reshape(3, n, 4)
    pub = _c(public_inputs).reshape(-1, 4)
    out = np.zeros(plonk_proof_words(lg_n), dtype=np.uint64)
    rc = _lib.orc_plonk_prove(curve, ctypes.c_uint(lg_n), _p(polys), _p(w_evals), _p(pub) if pub.shape[0] else None,
                              ctypes.c_size_t(pub.shape[0]), _p(_c(acc_prev)), _p(out))
    if rc != 0:
        raise OracleError(last_error())
    return out


def plonk_verify(curve, lg_n, C_qs, public_inputs, public_input_count, acc_prev, proof, succinct_only=False):
    pub = _c(public_inputs).reshape(-1, 4)
    rc = _lib.orc_plonk_verify(curve, ctypes.c_uint(lg_n), _p(_c(C_qs).reshape(10, 12)), _p(pub) if pub.shape[0] else None,
                               ctypes.c_size_t(pub.shape[0]), ctypes.c_size_t(public_input_count), _p(_c(acc_prev)),
                               _p(_c(proof)), int(succinct_only))
    if rc < 0:
        raise OracleError(last_error())
    return rc == 0


# ---------------- ark-serialize compressed encodings (oracle/serialize.hpp) ----------------
_u8p = ctypes.POINTER(ctypes.c_uint8)
_lib.orc_instance_serialize.restype = ctypes.c_size_t
_lib.orc_instance_deserialize.restype = ctypes.c_size_t


def point_compress(curve, xyz):
    out = (ctypes.c_uint8 * 33)()
    _lib.orc_point_compress(curve, _p(_c(xyz).reshape(12)), out)
    return bytes(out)


def point_decompress(curve, data):
    """Returns the Jacobian point, or None when the bytes are not a valid compressed point."""
    buf = (ctypes.c_uint8 * 33).from_buffer_copy(bytes(data[:33]))
    out = np.empty(12, dtype=np.uint64)
    return out if _lib.orc_point_decompress(curve, buf, _p(out)) == 0 else None


def instance_serialize(curve, inst):
    cap = 4096 + 66 * 64
    out = (ctypes.c_uint8 * cap)()
    n = _lib.orc_instance_serialize(curve, _p(_c(inst)), out, ctypes.c_size_t(cap))
    assert n > 0
    return bytes(out[:n])


def instance_deserialize(curve, data):
    """Returns (flat instance, bytes consumed) or None."""
    data = bytes(data)
    buf = (ctypes.c_uint8 * len(data)).from_buffer_copy(data)
    out = np.zeros(instance_words(40), dtype=np.uint64)
    lg = ctypes.c_uint(0)
    used = _lib.orc_instance_deserialize(curve, buf, ctypes.c_size_t(len(data)), _p(out), ctypes.c_size_t(out.shape[0]), ctypes.byref(lg))
    return (out[:instance_words(lg.value)].copy(), used) if used else None
