"""Constants of the Pasta cycle as the reference names them (crates/group/src/lib.rs:8-9)."""
import numpy as np

PALLAS, VESTA = 0, 1
FP, FQ = 0, 1   # halo Fp = ark_pallas::Fr (Pallas scalars), halo Fq = ark_pallas::Fq (Pallas base field)

MODULUS = {
    FP: 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001,
    FQ: 0x40000000000000000000000000000000224698fc094cf91b992d30ed00000001,
}
SCALAR_FIELD = {PALLAS: FP, VESTA: FQ}
BASE_FIELD = {PALLAS: FQ, VESTA: FP}
R = 1 << 256
TWO_ADICITY = 32
GENERATOR = 5


def int_to_limbs(x):
    return np.array([(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(4)], dtype=np.uint64)


def limbs_to_int(l):
    return sum(int(v) << (64 * i) for i, v in enumerate(l))


def to_mont(field, x):
    return int_to_limbs((x % MODULUS[field]) * R % MODULUS[field])


def from_mont(field, limbs):
    return limbs_to_int(limbs) * pow(R, -1, MODULUS[field]) % MODULUS[field]


def one(field):
    return to_mont(field, 1)


def affine_to_projective(curve, xy):
    """WrappedPoint -> Jacobian with Z = 1 (identity (0,0) -> (1,1,0)), like Affine::into()."""
    xy = np.asarray(xy, dtype=np.uint64).reshape(-1, 8)
    out = np.zeros((xy.shape[0], 12), dtype=np.uint64)
    o = one(BASE_FIELD[curve])
    for i in range(xy.shape[0]):
        if not xy[i].any():
            out[i, 0:4] = o
            out[i, 4:8] = o
        else:
            out[i, 0:8] = xy[i]
            out[i, 8:12] = o
    return out


def root_of_unity(field, log_n):
    """omega_N = (5^((p-1)/2^32))^(2^(32-log N)) as a canonical integer (ark-ff GENERATOR = 5, TWO_ADICITY = 32;
    pinned by crates/plonk/src/frontend/ivc/mod.rs:56,113)."""
    p = MODULUS[field]
    w = pow(GENERATOR, (p - 1) >> TWO_ADICITY, p)
    return pow(w, 1 << (TWO_ADICITY - log_n), p)


def root_power(field, log_n, i):
    """omega_N^i in Montgomery limbs (host-side scalar, not a hot path)."""
    return to_mont(field, pow(root_of_unity(field, log_n), i, MODULUS[field]))
