"""Oracle parity at the sizes BASELINE.json's configs name (the round-1 suite stopped at 2^10 rows / m = 5):

  config 4  PLONK proofs of 2^16-row circuits on both curves, field by field against the oracle's proof: the
            multiplication circuit of SURVEY section 8(d) (sparse columns, derived identity columns, staged upload,
            circuit-resident variant) and a circuit with EVERY selector, witness, round-constant and permutation column
            dense and random (unsatisfied: all gate expressions, the 16-coset quotient, no derived columns)
            ≙ crates/plonk/src/plonk/protocol.rs:64-502.  PARITY UNPINNED: the reference stores no proof, so the
            oracle's PLONK restatement is anchored only through the pinned primitives it calls.
  config 5  AS_DL accumulator equality at m = 64 instances of n = 2^12 and n = 2^16, accept / reject agreement with the
            oracle at m = 256 and m = 1024 with one tampered instance in the middle of the batch (the threaded
            succinct-check / sponge hand-off)  ≙ crates/accumulation/src/acc.rs:128-217, benches/acc.rs:28-83.
  a2        point_dot on 1024 Jacobian bases with identities, duplicates and a negated pair  ≙ group.rs:53-56, acc.rs:166.
  config 3  coset NTT / iNTT of 16 x 2^20 elements, bit for bit.
  a4        pcdl::chunked_commit (pcdl.rs:294-314) and HPoly::eval at lg n = 0 (pcdl.rs:221-233).
"""
import ctypes

import numpy as np
import pytest

import plonk_fixture as F
from test_pcdl_acc_gpu import same_instance
from test_plonk_gpu import same_proof

pytestmark = pytest.mark.gpu

LG = 16
N = 1 << LG


@pytest.fixture(scope="module")
def hb():
    import halo_b200
    halo_b200.init()
    return halo_b200


@pytest.fixture(scope="module")
def params16(hb, oracle):
    """The reference's own generator derivation (crates/group/src/main.rs:55-117) at 2^16, both curves."""
    out = {}
    for curve in (0, 1):
        gs, S, H = oracle.pp_generate(curve, N)
        hb.pp.reset(curve)
        hb.pp.set_pp(curve, gs, S, H)
        oracle.pp_set(curve, gs, S, H)
        out[curve] = (gs, S, H)
    yield out
    for curve in (0, 1):
        hb.pp.reset(curve)


# ------------------------------------------------------------------------------------------------------- config 4
@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("kind,n_public", [("mul", 3), ("random", 2)])
def test_plonk_2p16_proof_equals_oracle(hb, oracle, params16, curve, kind, n_public):
    O = oracle
    c = F.make_circuit(O, curve, LG, kind, seed=1600 + curve, n_public=n_public)
    acc0 = hb.acc.Accumulator.zero(curve, N)
    circuit = hb.plonk.PlonkCircuit.from_polys(curve, c["polys"], c["n_public"])
    pub = hb.plonk.PlonkPublicInputs(c["public_inputs"], acc0)
    wit = hb.plonk.PlonkWitness(c["polys"], c["w_evals"])
    pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    exp = O.plonk_prove(curve, LG, c["polys"], c["w_evals"], c["public_inputs"], acc0.q.w)
    same_proof(O, curve, LG, pi.w, exp)
    # circuit columns resident in HBM, only the witness uploaded: the identical proof
    circuit.make_resident(c["polys"])
    try:
        res = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    finally:
        circuit.release()
    same_proof(O, curve, LG, res.w, exp)
    # the witness commitments run beside the rest of round 0 and the two openings side by side (second execution
    # context); strictly one after the other (tuning keys 15 and 13) gives the identical proof
    from halo_b200._lib import lib
    lib.halo_set_tuning(15, 1); lib.halo_set_tuning(13, 1)
    try:
        seq = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    finally:
        lib.halo_set_tuning(15, 0); lib.halo_set_tuning(13, 0)
    same_proof(O, curve, LG, seq.w, exp)
    ok = O.plonk_verify(curve, LG, circuit.C_qs, c["public_inputs"], c["n_public"], acc0.q.w, pi.w)
    if kind == "random":
        assert not ok
        with pytest.raises(hb.Rejected):
            pi.verify_succinct(circuit, pub)
    else:
        assert ok
        pi.verify(circuit, pub)


# ------------------------------------------------------------------------------------------------------- config 5
def _instances(hb, O, curve, lg_n, m, seed):
    sf = O.SCALAR_FIELD[curve]
    n = 1 << lg_n
    qs = []
    for i in range(m):
        p = O.random_scalars(sf, n, seed + i)
        if i % 7 == 3:
            p[n // 2:] = 0                                # degree below the bound
        z = O.random_scalars(sf, 1, seed + 100000 + i)[0]
        qs.append(hb.pcdl.Instance.open(curve, p, n - 1, z))
    return qs


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("lg_n", [12, 16])
def test_accumulate_64_instances_equals_oracle(hb, oracle, params16, curve, lg_n):
    O = oracle
    m = 64
    qs = _instances(hb, O, curve, lg_n, m, 5000 * lg_n + curve)
    flat = np.stack([q.w for q in qs])
    assert all(O.pcdl_succinct_check(curve, q.w)[0] for q in qs[::9])      # spot check of the inputs by the oracle
    assert O.pcdl_check(curve, qs[5].w)
    acc = hb.acc.prover(curve, qs)
    exp = O.acc_prover(curve, flat, lg_n)
    same_instance(O, curve, acc.q.w, exp)
    hb.acc.verifier(curve, qs, acc)
    hb.acc.decider(curve, acc)
    assert O.acc_verifier(curve, flat, lg_n, acc.q.w) and O.acc_decider(curve, acc.q.w)


@pytest.mark.parametrize("curve,m", [(0, 256), (1, 1024)])
def test_large_batches_accept_and_reject_like_the_oracle(hb, oracle, params16, curve, m):
    O = oracle
    lg_n = 12
    sf = O.SCALAR_FIELD[curve]
    base = _instances(hb, O, curve, lg_n, 32, 7000 + curve)
    qs = [base[i % 32] for i in range(m)]                                   # repeated instances are legal inputs
    flat = np.stack([q.w for q in qs])
    acc = hb.acc.prover(curve, qs)
    hb.acc.verifier(curve, qs, acc)
    hb.acc.decider(curve, acc)
    assert O.acc_verifier(curve, flat, lg_n, acc.q.w) and O.acc_decider(curve, acc.q.w)
    if m <= 256:                                                            # accumulator equality where the oracle is quick
        same_instance(O, curve, acc.q.w, O.acc_prover(curve, flat, lg_n))
    # one tampered instance in the middle of the batch: v + 1
    bad = qs[m // 2].w.copy()
    bad[17:21] = O.field_add(sf, bad[17:21], O.ints_to_mont(sf, [1])[0])[0]
    qs_bad = list(qs)
    qs_bad[m // 2] = hb.pcdl.Instance(curve, bad)
    with pytest.raises(hb.Rejected):
        hb.acc.prover(curve, qs_bad)
    with pytest.raises(hb.Rejected):
        hb.acc.verifier(curve, qs_bad, acc)
    flat_bad = flat.copy()
    flat_bad[m // 2] = bad
    assert not O.acc_verifier(curve, flat_bad, lg_n, acc.q.w)
    # the untampered batch against an accumulator of another batch
    other = hb.acc.prover(curve, qs[:-1])
    with pytest.raises(hb.Rejected):
        hb.acc.verifier(curve, qs, other)
    assert not O.acc_verifier(curve, flat, lg_n, other.q.w)


@pytest.mark.parametrize("curve", [0, 1])
def test_malformed_instances_are_rejected_not_read_out_of_bounds(hb, oracle, params16, curve):
    """d of the first instance sizes every later read; non-canonical scalars and off-curve points."""
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    bf = 1 - sf
    qs = _instances(hb, O, curve, 6, 3, 9100 + curve)
    acc = hb.acc.prover(curve, qs)
    P = O.MODULUS
    for d_bad in (2 ** 40 - 1, 5, 2 ** 64 - 1, 2 ** 20 - 1):               # huge, not 2^k - 1, wrapping, above D
        w = qs[0].w.copy()
        w[12] = np.uint64(d_bad)
        flat = np.concatenate([w] + [q.w for q in qs[1:]])
        out = np.zeros_like(w)
        rc = hb._lib.lib.halo_acc_prover(curve, hb._lib.ptr(flat), ctypes.c_size_t(3), hb._lib.ptr(out))
        assert rc in (hb._lib.HALO_EREJECT, hb._lib.HALO_EINVAL), rc
        rc = hb._lib.lib.halo_acc_verifier(curve, hb._lib.ptr(flat), ctypes.c_size_t(3), hb._lib.ptr(acc.q.w))
        assert rc in (hb._lib.HALO_EREJECT, hb._lib.HALO_EINVAL), rc
    lg = 6
    c_off = 21 + 24 * lg + 12

    def add_modulus(words, field):                                          # x + p < 2^256: same residue, other encoding
        v = sum(int(words[i]) << (64 * i) for i in range(4)) + P[field]
        assert v < 1 << 256
        return np.array([(v >> (64 * i)) & (2 ** 64 - 1) for i in range(4)], dtype=np.uint64)

    for off in (c_off, 13, 17):                                             # c, z, v
        w = qs[0].w.copy()
        w[off:off + 4] = add_modulus(w[off:off + 4], sf)
        with pytest.raises(hb.Rejected):
            hb.pcdl.Instance(curve, w).check()
        with pytest.raises(hb.Rejected):
            hb.acc.prover(curve, [hb.pcdl.Instance(curve, w)] + qs[1:])
    w = qs[0].w.copy()                                                      # L_1 off the curve: Y + 1
    y = w[21 + 12 + 4:21 + 12 + 8]
    w[21 + 12 + 4:21 + 12 + 8] = O.field_add(bf, y, O.ints_to_mont(bf, [1])[0])[0]
    with pytest.raises(hb.Rejected):
        hb.pcdl.Instance(curve, w).check()
    w = qs[0].w.copy()                                                      # C with a non-canonical X
    w[0:4] = add_modulus(w[0:4], bf)
    with pytest.raises(hb.Rejected):
        hb.pcdl.Instance(curve, w).succinct_check()


# ------------------------------------------------------------------------------------------------------- a2
@pytest.mark.parametrize("curve", [0, 1])
def test_point_dot_1024_jacobian_bases(hb, oracle, curve):
    O = oracle
    sf, bf = O.SCALAR_FIELD[curve], 1 - O.SCALAR_FIELD[curve]
    k = 1024
    aff = O.points_arith(curve, k, k0=0xABCDE + curve)
    aff[10] = aff[3]                                                        # duplicates
    aff[11] = aff[3]
    aff[20, 0:4] = aff[4, 0:4]                                              # a negated pair: (x, -y)
    aff[20, 4:8] = O.field_sub(bf, np.zeros((1, 4), dtype=np.uint64), aff[4, 4:8].reshape(1, 4))[0]
    aff[[7, 500, 1023]] = 0                                                 # identities
    sc = O.random_scalars(sf, k, 777 + curve)
    sc[20] = sc[4]                                                          # s*P + s*(-P) cancels
    sc[30] = 0
    # random Jacobian representatives (x z^2, y z^3, z); identities as (X, Y, 0) with arbitrary X, Y
    z = O.random_scalars(bf, k, 888 + curve)
    z2 = O.field_mul(bf, z, z)
    jac = np.zeros((k, 12), dtype=np.uint64)
    jac[:, 0:4] = O.field_mul(bf, aff[:, 0:4], z2)
    jac[:, 4:8] = O.field_mul(bf, aff[:, 4:8], O.field_mul(bf, z2, z))
    jac[:, 8:12] = z
    for i in (7, 500, 1023):
        jac[i, 0:8] = z[i].tolist() * 2
        jac[i, 8:12] = 0
    got = hb.group.point_dot(curve, sc, jac)
    exp = O.msm(curve, aff, sc)
    assert (O.point_to_affine(curve, got.reshape(1, 12)) == O.point_to_affine(curve, exp.reshape(1, 12))).all()
    # zip semantics (group.rs:53-56 pairs min(len) items)
    got = hb.group.point_dot(curve, sc[:700], jac)
    exp = O.msm(curve, aff[:700], sc[:700])
    assert (O.point_to_affine(curve, got.reshape(1, 12)) == O.point_to_affine(curve, exp.reshape(1, 12))).all()


# ------------------------------------------------------------------------------------------------------- config 3
@pytest.mark.parametrize("field", [0, 1])
def test_coset_ntt_16_x_2p20_bit_exact(hb, oracle, field):
    O = oracle
    b, lg = 16, 20
    x = O.random_scalars(field, b << lg, 2020 + field).reshape(b, 1 << lg, 4)
    g = O.ints_to_mont(field, [5])[0]
    fx = hb.poly.ntt(field, x, coset=g)
    assert (fx == O.ntt(field, x, coset=g)).all()
    assert (hb.poly.ntt(field, fx, inverse=True, coset=g) == x).all()
    ix = hb.poly.ntt(field, x, inverse=True, coset=g)
    assert (ix == O.ntt(field, x, inverse=True, coset=g)).all()
    # plain domain at the same shape
    assert (hb.poly.ntt(field, x[:4]) == O.ntt(field, x[:4])).all()


# ------------------------------------------------------------------------------------------------------- a4 / a8
@pytest.mark.parametrize("curve", [0, 1])
def test_chunked_commit_and_hpoly_eval_edge(hb, oracle, params16, curve):
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    n = 1 << 12
    p = O.random_scalars(sf, n, 4242 + curve)
    p[3000:] = 0                                                            # stripped length 3000
    w = O.random_scalars(sf, 1, 4343)[0]
    aff = lambda x: O.point_to_affine(curve, np.asarray(x, dtype=np.uint64).reshape(-1, 12))
    for chunk, hide in ((1024, None), (700, w), (3000, None), (4096, w), (1, None)):
        coeffs = p if chunk > 1 else p[:37]
        got = hb.pcdl.chunked_commit(curve, coeffs, n - 1, hide, chunk)
        stripped = coeffs[:3000]
        parts = [stripped[i:i + chunk] for i in range(0, stripped.shape[0], chunk)]
        assert got.shape[0] == len(parts)
        exp = np.stack([O.pcdl_commit(curve, c, n - 1, hide) for c in parts])
        assert (aff(got) == aff(exp)).all(), (chunk, hide is not None)
    assert hb.pcdl.chunked_commit(curve, np.zeros((5, 4), dtype=np.uint64), n - 1, None, 2).shape[0] == 0
    with pytest.raises(hb.PreconditionError):
        hb.pcdl.chunked_commit(curve, p, n - 2, None, 16)                   # n not a power of two
    with pytest.raises(hb.PreconditionError):
        hb.pcdl.chunked_commit(curve, p, 1023, None, 2048)                  # a chunk longer than Gs[0..n]
    # HPoly::eval with a single challenge (lg n = 0) is 1 + xi_0 z in the reference (pcdl.rs:225)
    for lg in (0, 1, 2, 7):
        xis = O.random_scalars(sf, lg + 1, 99 + lg)
        z = O.random_scalars(sf, 1, 98)[0]
        assert (hb.pcdl.HPoly(sf, xis).eval(z) == O.hpoly_eval(sf, xis, z)).all(), lg
