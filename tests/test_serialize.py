"""ark-serialize compressed wire formats (Instance / EvalProof / Accumulator, crates/accumulation/src/pcdl.rs:26,159,
acc.rs:22; file layout of crates/accumulation/src/main.rs:59-91): the oracle's restatement and the product's C-ABI
routines against the BYTES OF THE REFERENCE's own qs.bin (tests/golden/qs_sample.json), against each other, and
their rejection of malformed input.  Host code only: no GPU."""
import ctypes
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def S():
    import halo_b200
    return halo_b200.serialize


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(HERE, "golden", "qs_sample.json")))


def _aff(O, curve, xyz):
    return O.point_to_affine(curve, np.asarray(xyz, dtype=np.uint64).reshape(-1, 12))


def test_reference_bytes_round_trip_bit_for_bit(S, oracle, golden):
    """Every Instance in the reference's qs.bin sample parses, its points are on the curve, and serialising the parsed
    object reproduces the reference's bytes exactly -- through the oracle and through the product."""
    O = oracle
    curve = 0
    assert golden["file_bytes"] == 39965
    for e in golden["entries"]:
        lg = e["n"].bit_length() - 1
        for key in ("instance", "acc_q"):
            raw = bytes.fromhex(e[key])
            assert len(raw) == 253 + 66 * lg                       # hiding instances
            got = O.instance_deserialize(curve, raw)
            assert got is not None and got[1] == len(raw)
            flat_o = got[0]
            assert O.instance_serialize(curve, flat_o) == raw
            q, used = S.instance_deserialize(curve, raw)
            assert used == len(raw) and q.d == e["n"] - 1 and q.pi.hiding
            assert S.instance_serialize(q) == raw
            # both parsers produce the same group elements and field elements
            fo, fp = O.inst_fields(flat_o), O.inst_fields(q.w)
            for k in ("z", "v", "c", "w_prime"):
                assert (fo[k] == fp[k]).all(), k
            for k in ("C", "U", "C_bar", "Ls", "Rs"):
                assert (_aff(O, curve, fo[k]) == _aff(O, curve, fp[k])).all(), k
            pts = np.concatenate([fp["C"].reshape(1, 12), fp["U"].reshape(1, 12), fp["C_bar"].reshape(1, 12), fp["Ls"], fp["Rs"]])
            aff = _aff(O, curve, pts)
            assert all(O.point_on_curve(curve, a) for a in aff)
        # the legacy trailer: Vec of 2 scalars, a point, a scalar -- the same primitive encodings
        piv = bytes.fromhex(e["legacy_pi_V"])
        assert int.from_bytes(piv[:8], "little") == 2
        for off in (8, 40, 105):
            s = S.scalar_deserialize(0, piv[off:off + 32])
            assert S.scalar_serialize(0, s) == piv[off:off + 32]
        U = S.point_decompress(curve, piv[72:105])
        assert S.point_compress(curve, U) == piv[72:105] and O.point_compress(curve, U) == piv[72:105]


@pytest.mark.parametrize("curve", [0, 1])
def test_points_and_scalars_product_equals_oracle(S, oracle, curve):
    O = oracle
    sf, bf = O.SCALAR_FIELD[curve], O.BASE_FIELD[curve]
    pts = O.points_arith(curve, 200, k0=0x5e11a + curve)
    jac = O.affine_to_jac(curve, pts)
    z = O.random_scalars(bf, 200, 9)
    z2 = O.field_mul(bf, z, z)
    jac[:, 0:4] = O.field_mul(bf, pts[:, 0:4], z2)                      # non-trivial Jacobian representatives
    jac[:, 4:8] = O.field_mul(bf, pts[:, 4:8], O.field_mul(bf, z2, z))
    jac[:, 8:12] = z
    seen = set()
    for i in range(200):
        b = S.point_compress(curve, jac[i])
        assert b == O.point_compress(curve, jac[i]) and len(b) == 33 and b[32] in (0x00, 0x80)
        seen.add(b[32])
        back = S.point_decompress(curve, b)
        assert (_aff(O, curve, back) == pts[i]).all()
        assert (_aff(O, curve, O.point_decompress(curve, b)) == pts[i]).all()
        neg = bytes(b[:32]) + bytes([b[32] ^ 0x80])                      # the other root: -P
        m = _aff(O, curve, S.point_decompress(curve, neg))[0]
        assert (m[0:4] == pts[i, 0:4]).all() and (O.field_add(bf, m[4:8], pts[i, 4:8]) == 0).all()
    assert seen == {0x00, 0x80}
    ident = np.zeros(12, dtype=np.uint64); ident[0:4] = 7; ident[4:8] = 9       # Z = 0: the identity, whatever X and Y
    b = S.point_compress(curve, ident)
    assert b == bytes(32) + b"\x40" == O.point_compress(curve, ident)
    assert not _aff(O, curve, S.point_decompress(curve, b)).any()
    sc = O.random_scalars(sf, 100, 3)
    for s in sc:
        b = S.scalar_serialize(sf, s)
        assert int.from_bytes(b, "little") == O.from_mont(sf, s)
        assert (S.scalar_deserialize(sf, b) == s).all()


@pytest.mark.parametrize("curve", [0, 1])
def test_malformed_input_is_rejected(S, oracle, curve):
    import halo_b200 as hb
    O = oracle
    sf, bf = O.SCALAR_FIELD[curve], O.BASE_FIELD[curve]
    p_b, p_s = O.MODULUS[bf], O.MODULUS[sf]
    le = lambda v: int(v).to_bytes(32, "little")
    for bad in (le(p_b) + b"\x00", le(p_b + 5) + b"\x80", le((1 << 256) - 1) + b"\x00", le(3) + b"\xc0"):
        with pytest.raises(hb.PreconditionError):
            S.point_decompress(curve, bad)
        assert O.point_decompress(curve, bad) is None
    # x with x^3 + 5 a non-residue: about half of all x; product and oracle must agree on every candidate
    n_bad = 0
    for x in range(1, 40):
        ok_o = O.point_decompress(curve, le(x) + b"\x00") is not None
        try:
            S.point_decompress(curve, le(x) + b"\x00"); ok_p = True
        except hb.PreconditionError:
            ok_p = False
        assert ok_o == ok_p, x
        n_bad += not ok_p
    assert 5 < n_bad < 35
    for bad in (le(p_s), le(p_s + 1), le((1 << 256) - 1)):
        with pytest.raises(hb.PreconditionError):
            S.scalar_deserialize(sf, bad)
    assert (S.scalar_deserialize(sf, le(p_s - 1)) == O.ints_to_mont(sf, [p_s - 1])[0]).all()
    # bits 0..5 of the flag byte are ignored, as ark-ff's from_u8_remove_flags does
    g = S.point_compress(curve, O.generator(curve))
    assert (_aff(O, curve, S.point_decompress(curve, g[:32] + bytes([g[32] | 0x15]))) == _aff(O, curve, O.generator(curve))).all()


def test_instance_container_errors(S, oracle, golden):
    import halo_b200 as hb
    raw = bytes.fromhex(golden["entries"][2]["instance"])                 # n = 16, L = 4
    q, _ = S.instance_deserialize(0, raw)
    for cut in (10, 40, 41, 100, len(raw) - 1):
        with pytest.raises(hb.PreconditionError):
            S.instance_deserialize(0, raw[:cut])
        assert oracle.instance_deserialize(0, raw[:cut]) is None
    bad = bytearray(raw); bad[33:41] = (14).to_bytes(8, "little")           # d + 1 not a power of two
    with pytest.raises(hb.PreconditionError):
        S.instance_deserialize(0, bytes(bad))
    bad = bytearray(raw); bad[105:113] = (3).to_bytes(8, "little")          # Ls shorter than lg(d + 1)
    with pytest.raises(hb.PreconditionError):
        S.instance_deserialize(0, bytes(bad))
    tag = 33 + 8 + 64 + 2 * (8 + 33 * 4) + 33 + 32                           # the Option tag of C_bar
    assert raw[tag] == 1
    bad = bytearray(raw); bad[tag] = 2
    with pytest.raises(hb.PreconditionError):
        S.instance_deserialize(0, bytes(bad))
    # a non-hiding instance is 65 bytes shorter and round-trips as well
    w = q.w.copy(); L = 4
    w[21 + 24 * L + 16:] = 0
    nh = hb.pcdl.Instance(0, w)
    b = S.instance_serialize(nh)
    assert len(b) == 188 + 66 * L and b == oracle.instance_serialize(0, w)
    back, used = S.instance_deserialize(0, b)
    assert used == len(b) and not back.pi.hiding and S.instance_serialize(back) == b
    # the file container: Vec<(usize, Instance, Accumulator)> in the CURRENT layout (no trailer)
    ents = [(e["n"], S.instance_deserialize(0, bytes.fromhex(e["instance"]))[0],
             S.accumulator_deserialize(0, bytes.fromhex(e["acc_q"]))[0]) for e in golden["entries"]]
    blob = S.qs_serialize(ents)
    again = S.qs_deserialize(0, blob)
    assert [a[0] for a in again] == [e["n"] for e in golden["entries"]] and S.qs_serialize(again) == blob
