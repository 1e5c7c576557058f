"""The CPU restatement of the PLONK prover/verifier (oracle/plonk.hpp ≙ crates/plonk/src/plonk/protocol.rs:64-502) is
self-consistent: proofs of satisfiable synthetic circuits (multiplication rows, copy constraints, public inputs, the
equality / range-check / Poseidon gates) verify, proofs of unsatisfied ones are rejected.  CPU only."""
import numpy as np
import pytest

import plonk_fixture as F


@pytest.fixture(scope="module")
def genesis(oracle):
    O = oracle
    out = {}
    lg = 5
    n = 1 << lg
    for curve in (0, 1):
        gs, s, h = O.pp_generate(curve, n)
        O.pp_set(curve, gs, s, h)
        zero = O.instance_open(curve, np.zeros((0, 4), dtype=np.uint64), n - 1, np.zeros(4, dtype=np.uint64))
        out[curve] = O.acc_prover(curve, zero.reshape(1, -1), lg)          # ≙ Accumulator::zero(n, 1)
    return lg, out


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("kind,n_public,accept", [("mul", 0, True), ("mul", 3, True), ("copy", 0, True),
                                                  ("gates", 0, True), ("random", 2, False)])
def test_prove_then_verify(oracle, genesis, curve, kind, n_public, accept):
    O = oracle
    lg, acc0 = genesis
    n = 1 << lg
    c = F.make_circuit(O, curve, lg, kind, seed=3, n_public=n_public)
    pi = O.plonk_prove(curve, lg, c["polys"], c["w_evals"], c["public_inputs"], acc0[curve])
    assert pi.shape[0] == O.plonk_proof_words(lg)
    Cq = np.stack([O.pcdl_commit(curve, c["polys"][i], n - 1) for i in range(10)])
    assert O.plonk_verify(curve, lg, Cq, c["public_inputs"], c["n_public"], acc0[curve], pi) is accept
    if accept:
        bad = pi.copy()
        bad[0] ^= 1                                                       # a witness evaluation
        assert not O.plonk_verify(curve, lg, Cq, c["public_inputs"], c["n_public"], acc0[curve], bad, succinct_only=True)
        assert not O.plonk_verify(curve, lg, Cq, c["public_inputs"], c["n_public"] + 1, acc0[curve], pi, succinct_only=True)
