"""GPU parity of the NTT / iNTT / coset transforms and the Evals wrapper against the CPU oracle (bit-exact),
plus the properties the reference's own tests check (protocol.rs:1037-1118: interpolate(evaluate(a) o evaluate(b))
= a o b) and the golden 2^16-th roots of unity of ivc/mod.rs:56,113."""
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


@pytest.mark.parametrize("field", [0, 1])
@pytest.mark.parametrize("log_n", [0, 1, 2, 3, 5, 8, 10, 11, 12, 13, 16, 18])
def test_ntt_matches_oracle(hb, oracle, field, log_n):
    O = oracle
    n = 1 << log_n
    x = O.random_scalars(field, n, 100 + log_n)
    fwd = hb.poly.ntt(field, x)
    assert (fwd == O.ntt(field, x)).all()
    inv = hb.poly.ntt(field, x, inverse=True)
    assert (inv == O.ntt(field, x, inverse=True)).all()
    assert (hb.poly.ntt(field, fwd, inverse=True) == x).all()
    if log_n <= 8:
        assert (fwd == O.dft_naive(field, x)).all()


@pytest.mark.parametrize("field", [0, 1])
def test_ntt_batched_and_coset(hb, oracle, field):
    O = oracle
    log_n, b = 12, 5
    n = 1 << log_n
    x = np.stack([O.random_scalars(field, n, 7 + i) for i in range(b)])
    x[3] = 0
    got = hb.poly.ntt(field, x)
    for i in range(b):
        assert (got[i] == O.ntt(field, x[i])).all()
    g = O.ints_to_mont(field, [5])[0]
    for lg in (6, 12, 14):
        y = O.random_scalars(field, 1 << lg, lg)
        c = hb.poly.ntt(field, y, coset=g)
        assert (c == O.ntt(field, y, coset=g)).all()
        assert (hb.poly.ntt(field, c, inverse=True, coset=g) == y).all()
        assert (hb.poly.ntt(field, y, inverse=True, coset=g) == O.ntt(field, y, inverse=True, coset=g)).all()


@pytest.mark.parametrize("log_n", [20, 22])
def test_ntt_large_round_trip_and_linearity(hb, oracle, log_n):
    """Full BASELINE sizes: bit-exact against the oracle at 2^20, size-independent properties at 2^22."""
    O = oracle
    n = 1 << log_n
    x, y = O.random_scalars(0, n, 1), O.random_scalars(0, n, 2)
    fx = hb.poly.ntt(0, x)
    if log_n <= 20:
        assert (fx == O.ntt(0, x)).all()
    assert (hb.poly.ntt(0, fx, inverse=True) == x).all()
    fy = hb.poly.ntt(0, y)
    fxy = hb.poly.ntt(0, O.field_add(0, x, y))
    assert (fxy == O.field_add(0, fx, fy)).all()


@pytest.mark.parametrize("curve,name", [(0, "pallas"), (1, "vesta")])
def test_root_of_unity_matches_golden(hb, oracle, golden_dir, curve, name):
    """The NTT of X (coefficient vector e_1) is (1, omega, omega^2, ...): element 1 is the golden root."""
    G = json.load(open(os.path.join(golden_dir, "ivc_genesis.json")))[name]
    field = oracle.SCALAR_FIELD[curve]
    n = 1 << 16
    x = np.zeros((n, 4), dtype=np.uint64)
    x[1] = hb.consts.one(field)
    ev = hb.poly.ntt(field, x)
    assert (ev[1] == np.array(G["omega_mont"], dtype=np.uint64)).all()
    assert (ev[0] == hb.consts.one(field)).all()
    assert (hb.poly.Domain(field, n).element(1) == ev[1]).all()


@pytest.mark.parametrize("field", [0, 1])
def test_evals_wrapper(hb, oracle, field):
    """poly.rs semantics: rotate-by-one constructor, shifts, evaluate with padding / folding, pointwise ops,
    and protocol.rs:1079 evals_mul: interpolate(evaluate(a) * evaluate(b)) = a * b."""
    O = oracle
    P = hb.poly
    n = 64
    dom, big = P.Domain(field, n), P.Domain(field, 4 * n)
    v = O.random_scalars(field, n, 3)
    e = P.Evals.from_vec_and_domain(v, dom)
    assert (e.evals[1:] == v[:-1]).all() and (e.evals[0] == v[-1]).all()
    assert (e.shift_left().evals == v).all()
    # evaluate: short (zero-padded) and long (reduced mod X^N - 1) coefficient vectors
    a = O.random_scalars(field, 20, 4)
    pad = np.zeros((n, 4), dtype=np.uint64); pad[:20] = a
    assert (P.Evals.from_poly_ref(a, dom).evals == O.ntt(field, pad)).all()
    long = O.random_scalars(field, 3 * n + 5, 5)
    fold = np.zeros((n, 4), dtype=np.uint64)
    for k in range(0, long.shape[0], n):
        chunk = long[k:k + n]
        fold[: chunk.shape[0]] = O.field_add(field, fold[: chunk.shape[0]], chunk)
    assert (P.Evals.from_poly_ref(long, dom).evals == O.ntt(field, fold)).all()
    assert (P.Evals.from_poly_ref(np.zeros((0, 4), dtype=np.uint64), dom).evals == 0).all()
    # interpolate strips trailing zeros
    assert (P.Evals.from_poly_ref(a, dom).interpolate() == a).all()
    # multiplicativity on a domain large enough for the product
    b = O.random_scalars(field, 30, 6)
    prod = (P.Evals.from_poly_ref(a, big) * P.Evals.from_poly_ref(b, big)).interpolate()
    exp = np.zeros((49, 4), dtype=np.uint64)
    for i in range(20):
        exp[i:i + 30] = O.field_add(field, exp[i:i + 30], O.field_mul(field, np.tile(a[i], (30, 1)), b))
    assert (prod == exp).all()
    s = O.random_scalars(field, 1, 9)[0]
    ea, eb = P.Evals.from_poly_ref(a, dom), P.Evals.from_poly_ref(b, dom)
    assert ((ea + eb).evals == O.field_add(field, ea.evals, eb.evals)).all()
    assert ((ea - eb).evals == O.field_sub(field, ea.evals, eb.evals)).all()
    assert (ea.scale(s).evals == O.field_mul(field, ea.evals, np.tile(s, (n, 1)))).all()
    assert (ea.add_scalar(s).sub_scalar(s).evals == ea.evals).all()
    assert (ea.shift_left_small_domain(P.Domain(field, 16)).evals == np.roll(ea.evals, -4, axis=0)).all()
    # divide_by_vanishing: Z_H is zero on H itself -> the reference panics; on a larger domain it divides
    with pytest.raises(hb.PreconditionError):
        ea.divide_by_vanishing(dom)


@pytest.mark.parametrize("field", [0, 1])
@pytest.mark.parametrize("la,lb", [(1, 1), (2, 3), (20, 30), (257, 255), (1000, 24), (4096, 4096)])
def test_poly_mul_matches_schoolbook(hb, oracle, field, la, lb):
    """`&a * &b` (protocol.rs:132-139,198-199): exact product coefficients."""
    O = oracle
    p = O.MODULUS[field]
    a, b = O.random_scalars(field, la, la), O.random_scalars(field, lb, 7 * lb)
    got = hb.poly.poly_mul(field, a, b)
    ai = [O.from_mont(field, x) for x in a]
    bi = [O.from_mont(field, x) for x in b]
    if la * lb <= 1 << 16:
        exp = [0] * (la + lb - 1)
        for i, x in enumerate(ai):
            for j, y in enumerate(bi):
                exp[i + j] = (exp[i + j] + x * y) % p
        assert (got == O.ints_to_mont(field, exp)).all()
    else:                                   # large: check the product at a random point instead
        z = 0x1234567890abcdef1234567
        ev = lambda c: sum(v * pow(z, k, p) for k, v in enumerate(c)) % p
        assert ev([O.from_mont(field, x) for x in got]) == ev(ai) * ev(bi) % p
    assert hb.poly.poly_mul(field, a, np.zeros((3, 4), dtype=np.uint64)).shape[0] == 0


@pytest.mark.parametrize("field", [0, 1])
def test_scalar_dot_and_construct_powers(hb, oracle, field):
    """group.rs:43-45 and :58-66."""
    O = oracle
    p = O.MODULUS[field]
    for n in (0, 1, 5, 1000, 70001):
        a, b = O.random_scalars(field, max(n, 1), 3)[:n], O.random_scalars(field, max(n, 1), 4)[:n]
        exp = sum(O.from_mont(field, x) * O.from_mont(field, y) for x, y in zip(a, b)) % p
        assert O.from_mont(field, hb.group.scalar_dot(field, a, b)) == exp
    z = O.random_scalars(field, 1, 9)[0]
    zi = O.from_mont(field, z)
    pw = hb.group.construct_powers(field, z, 1000)
    assert [O.from_mont(field, x) for x in pw[:40]] == [pow(zi, k, p) for k in range(40)]
    assert O.from_mont(field, pw[999]) == pow(zi, 999, p)
    assert hb.group.construct_powers(field, z, 0).shape == (0, 4)
