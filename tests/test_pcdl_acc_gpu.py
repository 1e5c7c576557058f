"""GPU parity of PC_DL and AS_DL against the CPU oracle, through the C ABI, plus the reference's golden genesis
accumulators (crates/plonk/src/frontend/ivc/mod.rs:196-292) and the round-trip tests of pcdl.rs:597-758 and
acc.rs:246-309 (open -> check accepts, U = commit(h), prove -> verify -> decide, tampering is rejected)."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_PP = 1 << 12


@pytest.fixture(scope="module")
def hb():
    import halo_b200
    halo_b200.init()
    return halo_b200


@pytest.fixture(scope="module")
def params(hb, oracle):
    """pp for both curves on the device and in the oracle."""
    out = {}
    for curve in (0, 1):
        gs, S, H = oracle.pp_generate(curve, N_PP)
        hb.pp.reset(curve)
        hb.pp.set_pp(curve, gs, S, H)
        oracle.pp_set(curve, gs, S, H)
        out[curve] = (gs, S, H)
    yield out
    for curve in (0, 1):
        hb.pp.reset(curve)


def same_instance(O, curve, got, exp):
    """Bit-exact comparison of two flat instances: points as canonical affine, scalars as limbs."""
    g, e = O.inst_fields(got), O.inst_fields(exp)
    assert g["d"] == e["d"] and g["hiding"] == e["hiding"]
    for k in ("z", "v", "c", "w_prime"):
        assert (g[k] == e[k]).all(), k
    for k in ("C", "U", "C_bar"):
        if k == "C_bar" and not g["hiding"]:
            continue
        assert (O.point_to_affine(curve, g[k].reshape(1, 12)) == O.point_to_affine(curve, e[k].reshape(1, 12))).all(), k
    for k in ("Ls", "Rs"):
        assert (O.point_to_affine(curve, g[k]) == O.point_to_affine(curve, e[k])).all(), k


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("lg_n", [1, 2, 3, 5, 9, 12])
@pytest.mark.parametrize("hiding", [False, True])
def test_open_matches_oracle_and_checks(hb, oracle, params, curve, lg_n, hiding):
    """pcdl.rs:690 test_check / :701 test_check_no_hiding, with the proof compared field by field to the oracle."""
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    n = 1 << lg_n
    d = n - 1
    p = O.random_scalars(sf, n, 10 * lg_n + curve)
    if lg_n == 5:
        p[20:] = 0                                    # degree below the bound
    z = O.random_scalars(sf, 1, 3)[0]
    w = hq = hw = None
    if hiding:
        w, hw = O.random_scalars(sf, 2, 4)
        hq = O.random_scalars(sf, d, 5) if d else np.zeros((0, 4), dtype=np.uint64)
    q = hb.pcdl.Instance.open(curve, p, d, z, w, hq, hw)
    exp = O.instance_open(curve, p, d, z, w, hq, hw)
    same_instance(O, curve, q.w, exp)
    assert O.pcdl_check(curve, q.w)                   # the oracle accepts the GPU proof
    q.check()                                         # and so does the GPU path (full check incl. the size-n MSM)
    h, U = q.succinct_check()
    ok, xis, U_o = O.pcdl_succinct_check(curve, exp)
    assert ok and (h.xis == xis).all()
    assert (h.eval(z) == O.hpoly_eval(sf, xis, z)).all()
    assert (h.get_poly() == O.hpoly_coeffs(sf, xis)).all()
    # pcdl.rs:627 test_u_check: U (the folded generator) = commit(h)
    assert (O.point_to_affine(curve, hb.pcdl.commit(curve, h.get_poly(), d).reshape(1, 12))
            == O.point_to_affine(curve, U.reshape(1, 12))).all()


@pytest.mark.parametrize("curve", [0, 1])
def test_tampered_proofs_are_rejected(hb, oracle, params, curve):
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    n = 64
    p = O.random_scalars(sf, n, 77)
    z = O.random_scalars(sf, 1, 78)[0]
    q = hb.pcdl.Instance.open(curve, p, n - 1, z)
    for pos in (17, 21 + 12 * 2, 21 + 24 * 6 + 12):        # v, L_2, c
        bad = q.w.copy()
        if pos == 17 or pos == 21 + 24 * 6 + 12:
            bad[pos:pos + 4] = O.field_add(sf, bad[pos:pos + 4], O.ints_to_mont(sf, [1])[0])[0]
        else:
            bad[pos:pos + 12] = q.w[21:33]                 # replace L_2 by L_0
        with pytest.raises(hb.Rejected):
            hb.pcdl.Instance(curve, bad).check()
        assert not O.pcdl_check(curve, bad)
    # a valid succinct check with the wrong U must still fail the full check (pcdl.rs:580)
    with pytest.raises(hb.PreconditionError):
        hb.pcdl.commit(curve, p, n - 2)                    # n not a power of two
    with pytest.raises(hb.PreconditionError):
        hb.pcdl.commit(curve, p, n // 2 - 1)               # degree above the bound


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("lg_n,m", [(2, 2), (3, 3), (6, 5), (10, 4)])
def test_accumulation_round_trip_and_parity(hb, oracle, params, curve, lg_n, m):
    """acc.rs:246 test_acc_scheme: prover -> verifier -> decider, accumulator equal to the oracle's."""
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    n = 1 << lg_n
    qs = []
    for i in range(m):
        p = O.random_scalars(sf, n, 100 * lg_n + i)
        z = O.random_scalars(sf, 1, 200 + i)[0]
        hide = i % 2 == 1
        w, hw = (O.random_scalars(sf, 2, 300 + i) if hide else (None, None))
        hq = O.random_scalars(sf, n - 1, 400 + i) if hide else None
        qs.append(hb.pcdl.Instance.open(curve, p, n - 1, z, w, hq, hw))
    acc = hb.acc.prover(curve, qs)
    exp = O.acc_prover(curve, np.stack([q.w for q in qs]), lg_n)
    same_instance(O, curve, acc.q.w, exp)
    hb.acc.verifier(curve, qs, acc)
    hb.acc.decider(curve, acc)
    assert O.acc_verifier(curve, np.stack([q.w for q in qs]), lg_n, acc.q.w) and O.acc_decider(curve, acc.q.w)
    # accumulate the accumulator again with fresh instances (the IVC shape: acc_prev, q_r, q_r_omega)
    acc2 = hb.acc.prover(curve, [acc.q] + qs[:2])
    hb.acc.verifier(curve, [acc.q] + qs[:2], acc2)
    hb.acc.decider(curve, acc2)
    # a verifier given different instances rejects
    with pytest.raises(hb.Rejected):
        hb.acc.verifier(curve, qs[:-1], acc)
    bad = qs[0].w.copy(); bad[17] ^= np.uint64(1)
    with pytest.raises(hb.Rejected):
        hb.acc.prover(curve, [hb.pcdl.Instance(curve, bad)] + qs[1:])


@pytest.mark.parametrize("curve,name", [(0, "pallas"), (1, "vesta")])
def test_genesis_accumulator_matches_reference_golden(hb, oracle, params, golden_dir, curve, name):
    """Accumulator::zero(65536, 1) on the reference's real generators reproduces the constants hard-coded at
    crates/plonk/src/frontend/ivc/mod.rs:196-292: C, z, v, all 16 (L, R) pairs, U and c."""
    O = oracle
    G = json.load(open(os.path.join(golden_dir, "ivc_genesis.json")))[name]
    n = G["n"]
    gs, S, H = O.pp_generate(curve, n)
    hb.pp.reset(curve)
    hb.pp.set_pp(curve, gs, S, H)
    try:
        acc = hb.acc.Accumulator.zero(curve, n)
        f = O.inst_fields(acc.q.w)
        u = lambda k: np.array(G[k], dtype=np.uint64)
        aff = lambda x: O.point_to_affine(curve, np.asarray(x, dtype=np.uint64).reshape(-1, 12))
        assert (f["z"] == u("z")).all() and (f["v"] == u("v")).all() and (f["c"] == u("c")).all()
        assert (aff(f["C"]) == aff(u("C"))).all()
        assert (aff(f["U"]) == aff(u("U"))).all()
        assert (aff(f["Ls"]) == aff(u("Ls"))).all()
        assert (aff(f["Rs"]) == aff(u("Rs"))).all()
        hb.acc.decider(curve, acc)
    finally:
        hb.pp.reset(curve)
        hb.pp.set_pp(curve, *params[curve])          # back to the module's small parameter set


@pytest.mark.parametrize("curve", [0, 1])
def test_fold_variants_give_the_same_proof(hb, oracle, params, curve):
    """The generator fold is run (a) as the reference writes it, gs_l + xi*gs_r by a 255-bit ladder (pcdl.rs:427-429),
    (b) with xi split as k1 + k2*lambda, and (c) with the fold-free tail starting at different lengths; every variant
    must give the oracle's proof bit for bit."""
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    n = 1 << 11
    p = O.random_scalars(sf, n, 901)
    z = O.random_scalars(sf, 1, 902)[0]
    exp = O.instance_open(curve, p, n - 1, z)
    lib = hb._lib.lib
    # (literal ladder?, log2 of the tail length or -1 = never / 0 = default, radix-4 folds off?)
    for ladder, tail, no_r4 in ((1, -1, 0), (0, -1, 1), (0, -1, 0), (0, 0, 0), (0, 5, 1), (0, 3, 0), (0, 4, 0), (1, 8, 0), (0, 1, 0)):
        lib.halo_set_tuning(2, ladder); lib.halo_set_tuning(3, tail); lib.halo_set_tuning(4, no_r4)
        try:
            q = hb.pcdl.Instance.open(curve, p, n - 1, z)
        finally:
            lib.halo_set_tuning(2, 0); lib.halo_set_tuning(3, 0); lib.halo_set_tuning(4, 0)
        same_instance(O, curve, q.w, exp)
    # (d) the first fold from the fixed-base table of 2^k * Gs[n/2 + j] (no doublings), followed by ladder folds,
    # radix-4 pairs or the tail; a second opening of another size rebuilds the table, a third reuses it
    for tail, no_r4, size in ((-1, 0, n), (-1, 1, n), (5, 0, n), (3, 1, n), (-1, 0, n // 4), (2, 0, n // 4)):
        lib.halo_set_tuning(0, 2); lib.halo_set_tuning(3, tail); lib.halo_set_tuning(4, no_r4)
        try:
            q = hb.pcdl.Instance.open(curve, p[:size], size - 1, z)
        finally:
            lib.halo_set_tuning(0, 0); lib.halo_set_tuning(3, 0); lib.halo_set_tuning(4, 0)
        same_instance(O, curve, q.w, exp if size == n else O.instance_open(curve, p[:size], size - 1, z))


def test_wire_format_of_gpu_produced_objects(hb, oracle, params):
    """tools/gen_qs.py's content at small sizes: hiding instances and their accumulators produced on the GPU, written as
    Vec<(usize, Instance, Accumulator)>::serialize_compressed (crates/accumulation/src/main.rs:59-91), read back, and
    the parsed objects pass check / decider on the GPU and in the oracle; bytes equal the oracle's encoding."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import gen_qs
    O = oracle
    ents = gen_qs.generate(0, 2, 9, seed=7, log=lambda *_: None)
    blob = hb.serialize.qs_serialize(ents)
    # a hiding instance is 253 + 66 lg n bytes; the accumulator's opening of h(X) is not hiding (acc.rs:190): 188 + 66 lg n
    assert len(blob) == 8 + sum(8 + (253 + 66 * lg) + (188 + 66 * lg) for lg in range(2, 10))
    back = hb.serialize.qs_deserialize(0, blob)
    assert len(back) == 8
    for (n, q, a), (n2, q2, a2) in zip(ents, back):
        assert n == n2 and q2.d == n - 1 and q2.pi.hiding and not a2.q.pi.hiding
        same_instance(O, 0, q2.w, q.w)
        same_instance(O, 0, a2.q.w, a.q.w)
        q2.check()
        hb.acc.verifier(0, [q2], a2)
        hb.acc.decider(0, a2)
        assert O.pcdl_check(0, q2.w) and O.acc_decider(0, a2.q.w)
        assert hb.serialize.instance_serialize(q) == O.instance_serialize(0, q.w)
