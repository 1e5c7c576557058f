"""Pins the CPU oracle to the reference (vectors G1-G4).

Fixtures under tests/golden/ were extracted from /root/reference by
tools/extract_reference_data.py (data only).  No GPU needed.
"""
import json
import os

import numpy as np
import pytest


def _load(golden_dir, name):
    return json.load(open(os.path.join(golden_dir, name)))


@pytest.mark.parametrize("field", [0, 1])
def test_field_ops_match_python_bigints(oracle, field):
    O = oracle
    p = O.MODULUS[field]
    a = O.random_scalars(field, 200, 11)
    b = O.random_scalars(field, 200, 12)
    mul, add, sub, inv = O.field_mul(field, a, b), O.field_add(field, a, b), O.field_sub(field, a, b), O.field_inv(field, a)
    for i in range(200):
        x, y = O.from_mont(field, a[i]), O.from_mont(field, b[i])
        assert O.from_mont(field, mul[i]) == x * y % p
        assert O.from_mont(field, add[i]) == (x + y) % p
        assert O.from_mont(field, sub[i]) == (x - y) % p
        assert O.from_mont(field, inv[i]) == pow(x, -1, p)
    # edge values
    edge = O.ints_to_mont(field, [0, 1, p - 1, p - 2, 2])
    sq = O.field_mul(field, edge, edge)
    for e, s in zip([0, 1, p - 1, p - 2, 2], sq):
        assert O.from_mont(field, s) == e * e % p


def test_roots_of_unity_G3(oracle, golden_dir):
    g = _load(golden_dir, "ivc_genesis.json")
    # crates/plonk/src/frontend/ivc/mod.rs:56 (Fp) and :113 (Fq)
    assert [int(x) for x in oracle.root_of_unity(0, 16)] == g["pallas"]["omega_mont"]
    assert [int(x) for x in oracle.root_of_unity(1, 16)] == g["vesta"]["omega_mont"]
    for f in (0, 1):
        w = oracle.from_mont(f, oracle.root_of_unity(f, 20))
        assert pow(w, 1 << 20, oracle.MODULUS[f]) == 1 and pow(w, 1 << 19, oracle.MODULUS[f]) != 1


def test_poseidon_kats_G4(oracle, golden_dir):
    O = oracle
    k = _load(golden_dir, "poseidon_kats.json")

    def h2m(f, h):
        return O.to_mont(f, int.from_bytes(bytes.fromhex(h), "little"))

    # inner_sponge.rs:315-322: PoseidonSponge::<PallasConfig> works over Pallas' base field = halo Fq
    for v in k["kimchi_pallas_base"]:
        inp = np.array([h2m(1, h) for h in v["input"]], dtype=np.uint64).reshape(-1, 4)
        assert O.from_mont(1, O.poseidon_hash(1, inp)) == int.from_bytes(bytes.fromhex(v["output"]), "little")
    # inner_sponge.rs:325-368
    for v in k["manual"]:
        f = 0 if v["sponge_curve"] == "vesta" else 1
        inp = np.array([h2m(f, h) for h in v["input"]], dtype=np.uint64).reshape(-1, 4)
        assert O.from_mont(f, O.poseidon_hash(f, inp)) == int.from_bytes(bytes.fromhex(v["output"]), "little")


@pytest.mark.parametrize("curve,name", [(0, "pallas"), (1, "vesta")])
def test_generators_match_shipped_blobs(oracle, golden_dir, curve, name):
    """SHA3 derivation (crates/group/src/main.rs:55-68,113-116) reproduces the .precompute blobs,
    including the `i + k + 2` duplicate structure."""
    s = _load(golden_dir, "pp_sample.json")[name]
    gs, S, H = oracle.pp_generate(curve, 2 * 16384)
    assert (gs[:64] == np.array(s["gs00_first64"], dtype=np.uint64)).all()
    assert (gs[16384:16388] == np.array(s["gs01_first4"], dtype=np.uint64)).all()
    assert (gs[16382:16384] == np.array(s["gs00_last2"], dtype=np.uint64)).all()
    assert (gs[16384] == gs[1]).all()
    assert (S == np.array(s["S"], dtype=np.uint64)).all()
    assert (H == np.array(s["H"], dtype=np.uint64)).all()
    assert oracle.point_on_curve(curve, gs[100]) and oracle.point_on_curve(curve, S)


def test_blob_decoder_against_reference_file(oracle, golden_dir):
    path = "/root/reference/crates/group/.precompute/pallas/gs-63.bin"
    if not os.path.exists(path):
        pytest.skip("reference checkout not present (GPU box)")
    s = _load(golden_dir, "pp_sample.json")["pallas"]
    words = oracle.bincode_decode(open(path, "rb").read(), 16384 * 8)
    assert words.shape[0] == 16384 * 8
    assert (words.reshape(-1, 8)[-2:] == np.array(s["gs63_last2"], dtype=np.uint64)).all()
    # block 63 entry k is gen(63 + k + 2)
    gs, _, _ = oracle.pp_generate(0, 16384)
    assert (words.reshape(-1, 8)[0] == gs[63]).all()


@pytest.mark.parametrize("curve,name", [(0, "pallas"), (1, "vesta")])
def test_genesis_accumulator_G1_G2(oracle, golden_dir, curve, name):
    """Accumulator::zero(65536, 1) (ivc/mod.rs:196-292): one size-2^16 commit MSM, a full IPA
    opening (32 MSMs + complete generator fold), the whole Poseidon transcript."""
    O = oracle
    G = _load(golden_dir, "ivc_genesis.json")[name]
    n, L = G["n"], 16
    gs, S, H = O.pp_generate(curve, n)
    O.pp_set(curve, gs, S, H)
    q = O.instance_open(curve, np.zeros((0, 4), dtype=np.uint64), n - 1, np.zeros(4, dtype=np.uint64))
    ok, _, _ = O.pcdl_succinct_check(curve, q)
    assert ok
    acc = O.acc_prover(curve, q.reshape(1, -1), L)
    f = O.inst_fields(acc)

    def aff(p):
        return O.point_to_affine(curve, np.array(p, dtype=np.uint64).reshape(12))[0]

    assert (aff(f["C"]) == aff(G["C"])).all()
    assert [int(x) for x in f["z"]] == G["z"]
    assert [int(x) for x in f["v"]] == G["v"]
    assert [int(x) for x in f["c"]] == G["c"]
    assert (aff(f["U"]) == aff(G["U"])).all()
    for i in range(L):
        assert (aff(f["Ls"][i]) == aff(G["Ls"][i])).all(), i
        assert (aff(f["Rs"][i]) == aff(G["Rs"][i])).all(), i
    assert O.acc_verifier(curve, q.reshape(1, -1), L, acc)
    assert O.acc_decider(curve, acc)
    # the stored golden accumulator itself is accepted
    gold = acc.copy()
    gold[0:12] = np.array(G["C"], dtype=np.uint64).reshape(12)
    assert O.acc_decider(curve, gold)
