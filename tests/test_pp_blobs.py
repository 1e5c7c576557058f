"""The reference's parameter files (crates/group/.precompute, bincode 2 varint; pp.rs:36-53) through the library's
decoder.  CPU part: the varint reader on crafted blocks, the golden sample re-encoded, and (only in a container that
has the reference checkout) the shipped gs-00.bin / sh.bin themselves.  GPU part: PublicParams::new from blobs."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/crates/group/.precompute"


def varint(v):
    if v < 251:
        return bytes([v])
    if v < 1 << 16:
        return b"\xfb" + v.to_bytes(2, "little")
    if v < 1 << 32:
        return b"\xfc" + v.to_bytes(4, "little")
    return b"\xfd" + v.to_bytes(8, "little")


def encode_block(points, with_length=True):
    out = varint(len(points)) if with_length else b""
    for p in points:
        for limb in p:
            out += varint(int(limb))
    return out


def golden():
    return json.load(open(os.path.join(HERE, "golden", "pp_sample.json")))


def test_varint_reader_all_widths_and_errors():
    import halo_b200
    from halo_b200 import pp, _lib
    pts = [[0, 250, 251, 65535, 65536, (1 << 32) - 1, 1 << 32, (1 << 64) - 1], [7] * 8, [1 << 40] * 8]
    blob = encode_block(pts)
    got = pp.decode_block(blob)
    assert got.tolist() == pts
    assert pp.decode_block(blob, max_points=2).tolist() == pts[:2]            # .take(m)
    assert pp.decode_block(encode_block(pts[:2], with_length=False), with_length=False).tolist() == pts[:2]
    for bad in (blob[:-1], blob[:5], b"\xfe" + blob[1:], b""):
        with pytest.raises(_lib.HaloError):
            pp.decode_block(bad)


def test_golden_sample_round_trips_through_the_decoder():
    from halo_b200 import pp
    g = golden()
    for curve in ("pallas", "vesta"):
        first = g[curve]["gs00_first64"]
        assert pp.decode_block(encode_block(first)).tolist() == first
        sh = pp.decode_block(encode_block([g[curve]["S"], g[curve]["H"]], with_length=False), with_length=False)
        assert sh.tolist() == [g[curve]["S"], g[curve]["H"]]


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present (GPU box)")
def test_shipped_blobs_decode_to_the_golden_sample():
    from halo_b200 import pp
    g = golden()
    for curve in ("pallas", "vesta"):
        b0 = pp.decode_block(open(f"{REF}/{curve}/gs-00.bin", "rb").read())
        assert b0.shape == (16384, 8)
        assert b0[:64].tolist() == g[curve]["gs00_first64"] and b0[-2:].tolist() == g[curve]["gs00_last2"]
        b1 = pp.decode_block(open(f"{REF}/{curve}/gs-01.bin", "rb").read(), max_points=4)
        assert b1.tolist() == g[curve]["gs01_first4"]
        sh = pp.decode_block(open(f"{REF}/{curve}/sh.bin", "rb").read(), with_length=False)
        assert sh.tolist() == [g[curve]["S"], g[curve]["H"]]


@pytest.mark.gpu
@pytest.mark.parametrize("curve,name", [(0, "pallas"), (1, "vesta")])
def test_public_params_new_from_blobs(oracle, curve, name):
    """PublicParams::new(n) over two blocks: the first n generators become resident (commit over them equals the MSM
    over the same points passed explicitly), S and H come from sh.bin, a short block set is a parameter error and an
    off-curve point is rejected."""
    import halo_b200 as hb
    from halo_b200 import pp, _lib
    hb.init()
    O = oracle
    g = golden()[name]
    pts = g["gs00_first64"]
    blocks = [encode_block(pts[:40]), encode_block(pts[40:])]
    sh = encode_block([g["S"], g["H"]], with_length=False)
    pp.reset(curve)
    try:
        with pytest.raises(_lib.ParamsError):
            pp.load_blobs(curve, blocks[:1], sh, 64)
        params = pp.load_blobs(curve, blocks, sh, 64)
        assert pp.pp_len(curve) == 64 and params.S.tolist() == g["S"] and params.H.tolist() == g["H"]
        sc = O.random_scalars(O.SCALAR_FIELD[curve], 64, 5)
        bases = np.array(pts, dtype=np.uint64)
        got = hb.pcdl.commit(curve, sc, 63)
        exp = O.msm(curve, bases, sc)
        assert (O.point_to_affine(curve, got.reshape(1, 12)) == O.point_to_affine(curve, exp.reshape(1, 12))).all()
        pp.reset(curve)
        bad = [list(p) for p in pts]
        bad[3][0] ^= 1
        with pytest.raises(_lib.HaloError):
            pp.load_blobs(curve, [encode_block(bad)], sh, 64)
    finally:
        pp.reset(curve)
