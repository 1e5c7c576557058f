"""GPU parity of the PLONK prover and verifier against the CPU restatement (oracle/plonk.hpp ≙
crates/plonk/src/plonk/protocol.rs:64-502), through the C ABI: every evaluation, commitment, opening proof and the
next accumulator of the proof are compared field by field; GPU proofs verify under both the GPU and the oracle
verifier; tampered proofs and unsatisfied circuits are rejected."""
import numpy as np
import pytest

import plonk_fixture as F

pytestmark = pytest.mark.gpu

N_PP = 1 << 11


@pytest.fixture(scope="module")
def hb():
    import halo_b200
    halo_b200.init()
    return halo_b200


@pytest.fixture(scope="module")
def params(hb, oracle):
    for curve in (0, 1):
        gs, S, H = oracle.pp_generate(curve, N_PP)
        hb.pp.reset(curve)
        hb.pp.set_pp(curve, gs, S, H)
        oracle.pp_set(curve, gs, S, H)
    yield
    for curve in (0, 1):
        hb.pp.reset(curve)


def proof_fields(O, lg_n, w):
    w = np.asarray(w, dtype=np.uint64)
    ew, iw = 33 + 24 * lg_n, 54 + 24 * lg_n
    o = 78 * 4
    f = {"vs": w[:o].reshape(78, 4), "Cs": w[o:o + 33 * 12].reshape(33, 12)}
    o += 33 * 12
    f["pi_r"], f["pi_r_omega"] = w[o:o + ew], w[o + ew:o + 2 * ew]
    f["acc_next"] = w[o + 2 * ew:]
    assert f["acc_next"].shape[0] == iw
    return f


def same_proof(O, curve, lg_n, got, exp):
    from test_pcdl_acc_gpu import same_instance
    g, e = proof_fields(O, lg_n, got), proof_fields(O, lg_n, exp)
    assert (g["vs"] == e["vs"]).all(), np.flatnonzero((g["vs"] != e["vs"]).any(axis=1))
    ga, ea = O.point_to_affine(curve, g["Cs"]), O.point_to_affine(curve, e["Cs"])
    assert (ga == ea).all(), np.flatnonzero((ga != ea).any(axis=1))
    head = np.zeros(21, dtype=np.uint64)
    head[12] = (1 << lg_n) - 1
    for k in ("pi_r", "pi_r_omega"):                       # compare the EvalProofs through the instance comparer
        same_instance(O, curve, np.concatenate([head, g[k]]), np.concatenate([head, e[k]]))
    same_instance(O, curve, g["acc_next"], e["acc_next"])


def genesis(hb, curve, n):
    return hb.acc.Accumulator.zero(curve, n)


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("lg_n,kind,n_public", [(1, "mul", 0), (2, "mul", 1), (3, "copy", 0), (5, "mul", 3), (5, "gates", 0),
                                                (5, "random", 2), (6, "copy", 0), (8, "gates", 0), (9, "random", 0),
                                                (10, "mul", 5)])
def test_proof_matches_oracle_and_verifies(hb, oracle, params, curve, lg_n, kind, n_public):
    O = oracle
    n = 1 << lg_n
    c = F.make_circuit(O, curve, lg_n, kind, seed=11 + lg_n, n_public=n_public)
    acc0 = genesis(hb, curve, n)
    circuit = hb.plonk.PlonkCircuit.from_polys(curve, c["polys"], c["n_public"])
    pub = hb.plonk.PlonkPublicInputs(c["public_inputs"], acc0)
    wit = hb.plonk.PlonkWitness(c["polys"], c["w_evals"])
    pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    exp = O.plonk_prove(curve, lg_n, c["polys"], c["w_evals"], c["public_inputs"], acc0.q.w)
    same_proof(O, curve, lg_n, pi.w, exp)
    ok_oracle = O.plonk_verify(curve, lg_n, circuit.C_qs, c["public_inputs"], c["n_public"], acc0.q.w, pi.w)
    if kind == "random":
        assert not ok_oracle
        with pytest.raises(hb.Rejected):
            pi.verify_succinct(circuit, pub)
        return
    assert ok_oracle
    pi.verify_succinct(circuit, pub)
    pi.verify(circuit, pub)
    # a satisfying witness takes the 8-coset quotient; the 16-coset coefficient division gives the same proof
    from halo_b200._lib import lib
    lib.halo_set_tuning(5, 1)
    try:
        slow = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    finally:
        lib.halo_set_tuning(5, 0)
    same_proof(O, curve, lg_n, slow.w, exp)
    # all-zero columns are skipped (identity commitments, no transforms); treating them as dense gives the same proof
    lib.halo_set_tuning(6, 1)
    try:
        dense = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    finally:
        lib.halo_set_tuning(6, 0)
    same_proof(O, curve, lg_n, dense.w, exp)
    # the two openings of round 5 run side by side (second execution context); one after the other gives the same proof
    lib.halo_set_tuning(13, 1)
    try:
        serial = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    finally:
        lib.halo_set_tuning(13, 0)
    same_proof(O, curve, lg_n, serial.w, exp)
    # (tuning key 15 = 1: the witness commitments after round 0 instead of beside it; they overlap only from 2^12 rows)
    # circuit columns resident on the device (halo_plonk_circuit_load): only the witness is uploaded, same proof
    circuit.make_resident(c["polys"])
    try:
        res = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    finally:
        circuit.release()
    same_proof(O, curve, lg_n, res.w, exp)
    # the next accumulator is a valid one (protocol.rs:493-502) and chains into a second proof (IVC-style)
    hb.acc.decider(curve, pi.acc_next)
    pub2 = hb.plonk.PlonkPublicInputs(c["public_inputs"], pi.acc_next)
    pi2 = hb.plonk.PlonkProof.naive_prover(circuit, pub2, wit)
    pi2.verify(circuit, pub2)
    exp2 = O.plonk_prove(curve, lg_n, c["polys"], c["w_evals"], c["public_inputs"], pi.acc_next.q.w)
    same_proof(O, curve, lg_n, pi2.w, exp2)


@pytest.mark.parametrize("curve", [0, 1])
def test_tampering_is_rejected(hb, oracle, params, curve):
    O = oracle
    lg_n = 4
    n = 1 << lg_n
    c = F.make_circuit(O, curve, lg_n, "gates", seed=5)
    acc0 = genesis(hb, curve, n)
    circuit = hb.plonk.PlonkCircuit.from_polys(curve, c["polys"], 0)
    pub = hb.plonk.PlonkPublicInputs(c["public_inputs"], acc0)
    pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, hb.plonk.PlonkWitness(c["polys"], c["w_evals"]))
    pi.verify(circuit, pub)
    one = O.to_mont(c["field"], 1)
    for word in (0, 16 * 4, 41 * 4, 73 * 4, 74 * 4):       # a witness, round-constant, quotient, z, z_omega evaluation
        bad = pi.w.copy()
        bad[word:word + 4] = O.field_add(c["field"], bad[word:word + 4].reshape(1, 4), one.reshape(1, 4))[0]
        with pytest.raises(hb.Rejected):
            hb.plonk.PlonkProof(curve, lg_n, bad).verify_succinct(circuit, pub)
    bad = pi.w.copy()                                      # swap two witness commitments
    o = 78 * 4
    bad[o:o + 12], bad[o + 12:o + 24] = pi.w[o + 12:o + 24], pi.w[o:o + 12]
    with pytest.raises(hb.Rejected):
        hb.plonk.PlonkProof(curve, lg_n, bad).verify_succinct(circuit, pub)
    with pytest.raises(hb.Rejected):                       # wrong public-input count (protocol.rs:369)
        pi.verify_succinct(hb.plonk.PlonkCircuit(curve, n, 1, circuit.C_qs), pub)
    with pytest.raises(hb.Rejected):                       # another previous accumulator
        other = hb.acc.prover(curve, [hb.pcdl.Instance.zero(curve, n), hb.pcdl.Instance.zero(curve, n)])
        pi.verify_succinct(circuit, hb.plonk.PlonkPublicInputs(c["public_inputs"], other))


def test_zero_columns_and_bad_row_count(hb, oracle, params):
    """All-zero witness / permutation columns (every factor of f' and g' equals gamma) still prove bit-identically;
    a row count that is not a power of two is a precondition error (Domain::new / commit asserts)."""
    O = oracle
    curve, lg_n = 0, 3
    n = 1 << lg_n
    c = F.make_circuit(O, curve, lg_n, "mul", seed=2)
    polys = c["polys"].copy()
    polys[10:18] = 0
    polys[41:57] = 0
    acc0 = genesis(hb, curve, n)
    circuit = hb.plonk.PlonkCircuit.from_polys(curve, polys, 0)
    pub = hb.plonk.PlonkPublicInputs(c["public_inputs"], acc0)
    w_evals = np.zeros_like(c["w_evals"])
    pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, hb.plonk.PlonkWitness(polys, w_evals))
    exp = O.plonk_prove(curve, lg_n, polys, w_evals, c["public_inputs"], acc0.q.w)
    same_proof(O, curve, lg_n, pi.w, exp)
    # identity columns that do NOT follow id_i = id_0 + i*n (two of them swapped): the prover must notice and transform
    # all eight instead of deriving them from id_0; the proof still equals the oracle's on the same columns
    polys2 = c["polys"].copy()
    polys2[[42, 43]] = polys2[[43, 42]]
    pi2 = hb.plonk.PlonkProof.naive_prover(circuit, pub, hb.plonk.PlonkWitness(polys2, c["w_evals"]))
    exp2 = O.plonk_prove(curve, lg_n, polys2, c["w_evals"], c["public_inputs"], acc0.q.w)
    same_proof(O, curve, lg_n, pi2.w, exp2)
    with pytest.raises(hb.PreconditionError):
        hb.plonk.PlonkProof.naive_prover(hb.plonk.PlonkCircuit(curve, 3, 0, circuit.C_qs), pub, hb.plonk.PlonkWitness(polys, w_evals))


@pytest.mark.parametrize("curve", [0, 1])
def test_trace_new_matches_the_fixture_and_proves(hb, oracle, params, curve):
    """Trace::new / consume (trace.rs:137-283) from raw columns and copy-constraint classes: the polynomials equal the
    ones the fixture interpolates with the oracle, the 41 circuit commitments equal the oracle's commits, and the
    consumed triple proves and verifies."""
    O = oracle
    lg_n = 6
    n = 1 << lg_n
    c = F.make_circuit(O, curve, lg_n, "copy", seed=21)
    field = c["field"]
    # recover the column vectors from the fixture's polynomials (evaluate, undo the Evals rotation)
    cols = np.roll(O.ntt(field, c["polys"].copy()).reshape(57, n, 4), -1, axis=1)
    classes = [[(j + 2, 1), (j + 1, 3)] for j in range(0, n - 1, 2)]          # row j+1's a tied to row j's c (1-based)
    acc0 = genesis(hb, curve, n)
    tr = hb.plonk.Trace(curve, classes, c["public_inputs"], cols[10:26], cols[26:41], cols[0:10], n, acc0)
    assert (tr.polys == c["polys"]).all()
    assert (tr.w_evals[:3] == c["w_evals"]).all()
    exp_C = np.stack([O.pcdl_commit(curve, c["polys"][i], n - 1) for i in list(range(10)) + list(range(26, 57))])
    got_C = np.concatenate([tr.C_qs, tr.C_rs, tr.C_ids, tr.C_sigmas])
    assert (O.point_to_affine(curve, got_C) == O.point_to_affine(curve, exp_C)).all()
    circuit, pub, wit = tr.consume()
    pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, wit)
    pi.verify(circuit, pub)
    assert O.plonk_verify(curve, lg_n, circuit.C_qs, c["public_inputs"], 0, acc0.q.w, pi.w)
    with pytest.raises(hb.PreconditionError):
        hb.plonk.Trace(curve, [], c["public_inputs"], cols[10:26, :48], cols[26:41, :48], cols[0:10, :48], 48, acc0)


def test_large_proof_verifies_and_matches_commit_path(hb, oracle):
    """2^15 rows (60 MB of columns: the staged multi-threaded upload path) on synthetic generators: the proof verifies,
    its witness commitments equal stand-alone commits of the same columns, and a different witness is rejected
    against it.  Size-independent properties only - the oracle prover is not run at this size."""
    import ctypes
    from halo_b200._lib import lib, check
    curve, lg_n = 0, 15
    n = 1 << lg_n
    u64p = ctypes.POINTER(ctypes.c_uint64)
    gs = np.zeros((n + 2, 8), dtype=np.uint64)
    check(lib.halo_gen_points(curve, ctypes.c_uint64(77), ctypes.c_size_t(0), ctypes.c_size_t(n + 2), gs.ctypes.data_as(u64p)))
    hb.pp.reset(curve)
    hb.pp.set_pp(curve, gs[:n], gs[n], gs[n + 1])
    try:
        polys, w_evals = hb.plonk.synthetic_circuit(curve, lg_n, seed=3)
        acc0 = hb.acc.Accumulator.zero(curve, n)
        circuit = hb.plonk.PlonkCircuit.from_polys(curve, polys, 0)
        pub = hb.plonk.PlonkPublicInputs(np.zeros((0, 4), dtype=np.uint64), acc0)
        pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, hb.plonk.PlonkWitness(polys, w_evals))
        pi.verify(circuit, pub)
        C_ws = pi.commitments()["ws"]
        for i in (0, 1, 2):
            exp = hb.pcdl.commit(curve, polys[10 + i], n - 1)
            assert (oracle.point_to_affine(curve, C_ws[i].reshape(1, 12)) == oracle.point_to_affine(curve, exp.reshape(1, 12))).all()
        # z = 1 for the identity permutation: its commitment is G_0 and its evaluations are one
        one = hb.consts.one(curve)
        ev = pi.evals()
        assert (ev["z"] == one).all() and (ev["z_omega"] == one).all()
        bad = pi.w.copy()
        bad[4:8] = ev["ws"][0]                              # w_1(xi) := w_0(xi)
        with pytest.raises(hb.Rejected):
            hb.plonk.PlonkProof(curve, lg_n, bad).verify_succinct(circuit, pub)
    finally:
        hb.pp.reset(curve)


@pytest.mark.parametrize("curve", [0, 1])
def test_all_gates_synthetic_circuit_satisfied(hb, oracle, curve):
    """The benchmark's all-gates circuit (every selector, witness and round-constant column random) with its constant
    selector solved row by row from halo_plonk_gate_residual: f_GC vanishes on H, the proof equals the oracle's field by
    field, and BOTH verifiers accept it -- every custom gate expression is live in an accepted proof.  The unsolved
    circuit has a non-zero residual and is rejected."""
    O = oracle
    lg_n = 8
    n = 1 << lg_n
    none = np.zeros((0, 4), dtype=np.uint64)
    gs, S, H = O.pp_generate(curve, N_PP)                 # (an earlier test of this module resets the parameters)
    hb.pp.reset(curve); hb.pp.set_pp(curve, gs, S, H); O.pp_set(curve, gs, S, H)
    polys, w_evals = hb.plonk.synthetic_circuit(curve, lg_n, seed=21 + curve, dense_gates=True, satisfied=True)
    assert not hb.plonk.gate_residual(curve, polys, w_evals).any()
    acc0 = genesis(hb, curve, n)
    circuit = hb.plonk.PlonkCircuit.from_polys(curve, polys, 0)
    pub = hb.plonk.PlonkPublicInputs(none, acc0)
    pi = hb.plonk.PlonkProof.naive_prover(circuit, pub, hb.plonk.PlonkWitness(polys, w_evals))
    same_proof(O, curve, lg_n, pi.w, O.plonk_prove(curve, lg_n, polys, w_evals, none, acc0.q.w))
    pi.verify(circuit, pub)
    assert O.plonk_verify(curve, lg_n, circuit.C_qs, none, 0, acc0.q.w, pi.w)
    bad_polys, bad_evals = hb.plonk.synthetic_circuit(curve, lg_n, seed=21 + curve, dense_gates=True)
    assert hb.plonk.gate_residual(curve, bad_polys, bad_evals).any()
    bad_circuit = hb.plonk.PlonkCircuit.from_polys(curve, bad_polys, 0)
    bad = hb.plonk.PlonkProof.naive_prover(bad_circuit, pub, hb.plonk.PlonkWitness(bad_polys, bad_evals))
    with pytest.raises(hb.Rejected):
        bad.verify_succinct(bad_circuit, pub)
    hb.pp.reset(curve)
