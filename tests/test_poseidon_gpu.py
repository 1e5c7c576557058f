"""Poseidon on the device (halo_poseidon_hash_batch) against the oracle sponge, which is
pinned on the Kimchi known-answer vectors (crates/poseidon/test-vectors/kimchi-vecs.json, inner_sponge.rs:325-368): the
KAT inputs themselves, every input length 0..9 (both rate cells, the empty sponge), edge values, and a large batch."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hb():
    import halo_b200
    halo_b200.init()
    return halo_b200


def test_kimchi_known_answers_on_the_device(hb, oracle, golden_dir):
    O = oracle
    k = json.load(open(os.path.join(golden_dir, "poseidon_kats.json")))
    h2m = lambda f, h: O.to_mont(f, int.from_bytes(bytes.fromhex(h), "little"))
    cases = [(1, v) for v in k["kimchi_pallas_base"]] + [(0 if v["sponge_curve"] == "vesta" else 1, v) for v in k["manual"]]
    for f, v in cases:
        inp = np.array([h2m(f, h) for h in v["input"]], dtype=np.uint64).reshape(1, -1, 4)
        got = hb.poseidon.hash_batch(f, inp)[0]
        assert O.from_mont(f, got) == int.from_bytes(bytes.fromhex(v["output"]), "little")


@pytest.mark.parametrize("field", [0, 1])
def test_batches_equal_the_oracle_sponge(hb, oracle, field):
    O = oracle
    p = O.MODULUS[field]
    for length in range(0, 10):
        count = 37
        x = O.random_scalars(field, count * max(1, length), 100 + length).reshape(count, -1, 4)[:, :length]
        if length:
            edge = O.ints_to_mont(field, [0, 1, p - 1, p - 2])
            x[0, 0] = edge[0]; x[1, -1] = edge[2]; x[2, 0] = edge[1]; x[3, :] = edge[3]
        got = hb.poseidon.hash_batch(field, np.ascontiguousarray(x))
        for i in range(count):
            assert (got[i] == O.poseidon_hash(field, x[i])).all(), (length, i)
    # a large batch (the size where the device pays): 2^14 sponges of two elements, spot-checked
    big = O.random_scalars(field, 2 << 14, 7).reshape(1 << 14, 2, 4)
    got = hb.poseidon.hash_batch(field, big)
    for i in (0, 1, 4095, (1 << 14) - 1):
        assert (got[i] == O.poseidon_hash(field, big[i])).all()
    assert hb.poseidon.hash_batch(field, np.zeros((0, 3, 4), dtype=np.uint64)).shape == (0, 4)
