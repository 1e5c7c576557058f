"""GPU parity of the field arithmetic and the MSM path against the CPU oracle, through the C ABI.
Mirrors what the reference's tests check for this path (pedersen.rs:68 homomorphism, pp.rs:126 on-curve params)
plus bit-exact known answers from the oracle and the golden commitment of ivc/mod.rs:197."""
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


def _aff(O, curve, xyz):
    return O.point_to_affine(curve, np.asarray(xyz, dtype=np.uint64).reshape(-1, 12))


@pytest.mark.parametrize("field", [0, 1])
def test_field_ops_bit_exact(hb, oracle, field):
    O = oracle
    p = O.MODULUS[field]
    n = 20000
    a, b = O.random_scalars(field, n, 5), O.random_scalars(field, n, 6)
    edge = O.ints_to_mont(field, [0, 1, p - 1, p - 2, 2, (1 << 254) - 1, 1 << 32, (1 << 32) - 1, p >> 1])
    a[: len(edge)] = edge
    b[3: 3 + len(edge)] = edge
    assert (hb.group.field_op(field, "mul", a, b) == O.field_mul(field, a, b)).all()
    assert (hb.group.field_op(field, "add", a, b) == O.field_add(field, a, b)).all()
    assert (hb.group.field_op(field, "sub", a, b) == O.field_sub(field, a, b)).all()
    assert (hb.group.field_op(field, "inv", a[:200]) == O.field_inv(field, a[:200])).all()
    c = hb.group.field_op(field, "to_canonical", a)
    assert (c == O.field_to_canonical(field, a)).all()
    assert (hb.group.field_op(field, "from_canonical", c) == a).all()


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("k", [0, 1, 2, 3, 7, 33, 100, 1 << 8, 1000, 1 << 12, (1 << 14) + 5, 1 << 16])
def test_msm_random_points(hb, oracle, curve, k):
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    pts = O.points_arith(curve, max(k, 1))[:k]
    sc = O.random_scalars(sf, max(k, 1), 1000 + k)[:k]
    got = hb.group.point_dot_affine(curve, sc, pts)
    exp = O.msm(curve, pts, sc) if k else O.affine_to_jac(curve, np.zeros(8, dtype=np.uint64))[0]
    assert (_aff(O, curve, got) == _aff(O, curve, exp)).all()


@pytest.mark.parametrize("curve", [0, 1])
def test_msm_edge_scalars_and_identity_bases(hb, oracle, curve):
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    k = 5000
    pts = O.points_arith(curve, k).copy()
    pts[17] = 0            # identity base
    pts[18] = pts[19]      # duplicate base
    pts[21] = pts[20]; pts[21, 4:8] = O.field_sub(O.BASE_FIELD[curve], np.zeros(4, dtype=np.uint64), pts[20, 4:8])[0]  # -P
    sc = O.edge_scalars(sf, k, 99)
    sc[20] = sc[21]        # s*P + s*(-P)
    got = hb.group.point_dot_affine(curve, sc, pts)
    assert (_aff(O, curve, got) == _aff(O, curve, O.msm(curve, pts, sc))).all()
    # all-zero scalars -> identity
    z = hb.group.point_dot_affine(curve, np.zeros((64, 4), dtype=np.uint64), pts[:64])
    assert not _aff(O, curve, z).any()
    # unchecked semantics: min(len) pairs
    got = hb.group.point_dot_affine(curve, sc[:100], pts[:60])
    assert (_aff(O, curve, got) == _aff(O, curve, O.msm(curve, pts[:60], sc[:60]))).all()


@pytest.mark.parametrize("c", [4, 7, 11, 13, 16])
def test_msm_every_window_width(hb, oracle, c):
    import ctypes
    O = oracle
    k = 3000
    pts, sc = O.points_arith(0, k), O.random_scalars(0, k, c)
    hb._lib.lib.halo_msm_set_window(ctypes.c_int(c))
    try:
        got = hb.group.point_dot_affine(0, sc, pts)
    finally:
        hb._lib.lib.halo_msm_set_window(ctypes.c_int(0))
    assert (_aff(O, 0, got) == _aff(O, 0, O.msm(0, pts, sc))).all()


@pytest.mark.parametrize("curve,name", [(0, "pallas"), (1, "vesta")])
def test_resident_generators_commit_and_golden_C(hb, oracle, golden_dir, curve, name):
    """pcdl::commit path on the reference's real (duplicate-heavy) generators; reproduces the golden
    acc_0.q.C = commit(h_0) of crates/plonk/src/frontend/ivc/mod.rs:197,246."""
    O = oracle
    G = json.load(open(os.path.join(golden_dir, "ivc_genesis.json")))[name]
    n = G["n"]
    gs, S, H = O.pp_generate(curve, n)
    hb.pp.reset(curve)
    hb.pp.set_pp(curve, gs, S, H)
    assert hb.pp.pp_len(curve) == n
    with pytest.raises(hb.ParamsError):
        hb.pp.set_pp(curve, np.concatenate([gs, gs]), S, H)      # monotone like set_pp (pp.rs:63-82)
    hb.pp.set_pp(curve, gs[:1024], S, H)                         # smaller: no-op
    sf = O.SCALAR_FIELD[curve]
    sc = O.random_scalars(sf, n, 31337)
    got = hb.pedersen.commit(curve, None, (0, n), sc)
    assert (_aff(O, curve, got) == _aff(O, curve, O.msm(curve, gs, sc))).all()
    # hiding commit: w*S + MSM
    w = O.random_scalars(sf, 1, 7)[0]
    O.pp_set(curve, gs, S, H)
    got = hb.pedersen.commit(curve, w, (0, 4096), sc[:4096])
    assert (_aff(O, curve, got) == _aff(O, curve, O.pcdl_commit(curve, sc[:4096], 4095, w))).all()
    # golden: h_0 from the zero instance's transcript, C = commit(h_0)
    q = O.instance_open(curve, np.zeros((0, 4), dtype=np.uint64), n - 1, np.zeros(4, dtype=np.uint64))
    ok, xis, _ = O.pcdl_succinct_check(curve, q)
    assert ok
    h = O.hpoly_coeffs(sf, xis)
    got = hb.pedersen.commit(curve, None, (0, n), h)
    assert (_aff(O, curve, got) == _aff(O, curve, np.array(G["C"], dtype=np.uint64).reshape(12))).all()
    with pytest.raises(hb.PreconditionError):
        hb.pedersen.commit(curve, None, (0, 16), sc[:17])
    hb.pp.reset(curve)


def test_pp_rejects_off_curve_point(hb, oracle):
    gs, S, H = oracle.pp_generate(0, 64)
    bad = gs.copy(); bad[5, 0] ^= np.uint64(1)
    hb.pp.reset(0)
    with pytest.raises(hb.PreconditionError):
        hb.pp.set_pp(0, bad, S, H)
    assert hb.pp.pp_len(0) == 0


def test_pedersen_homomorphism(hb, oracle):
    """pedersen.rs:68 test_homomorphism_property."""
    O = oracle
    gs, S, H = O.pp_generate(0, 64)
    hb.pp.reset(0); hb.pp.set_pp(0, gs, S, H)
    m1, m2 = O.random_scalars(0, 2, 1), O.random_scalars(0, 2, 2)
    w1, w2 = O.random_scalars(0, 1, 3)[0], O.random_scalars(0, 1, 4)[0]
    inner = hb.pedersen.commit(0, O.field_add(0, w1, w2)[0], (0, 2), O.field_add(0, m1, m2))
    outer = O.point_add(0, hb.pedersen.commit(0, w1, (0, 2), m1), hb.pedersen.commit(0, w2, (0, 2), m2))
    assert (_aff(O, 0, inner) == _aff(O, 0, outer)).all()
    hb.pp.reset(0)


@pytest.mark.parametrize("k", [1 << 18, 1 << 20])
def test_msm_large_bit_exact(hb, oracle, k):
    O = oracle
    pts = O.points_arith(0, k)
    sc = O.random_scalars(0, k, k)
    got = hb.group.point_dot_affine(0, sc, pts)
    assert (_aff(O, 0, got) == _aff(O, 0, O.msm(0, pts, sc))).all()
    # linearity at full size: MSM(a) + MSM(b) == MSM(a+b)
    sc2 = O.random_scalars(0, k, k + 1)
    got2 = hb.group.point_dot_affine(0, sc2, pts)
    got3 = hb.group.point_dot_affine(0, O.field_add(0, sc, sc2), pts)
    assert (_aff(O, 0, got3) == _aff(O, 0, O.point_add(0, got, got2))).all()


@pytest.mark.parametrize("curve", [0, 1])
def test_msm_batch_over_resident_generators(hb, oracle, curve):
    """The 16-commit shape of a prover round (protocol.rs:114,263): one pipeline, same bases, b scalar vectors."""
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    n, b = 4096, 7
    gs, S, H = O.pp_generate(curve, n)
    hb.pp.reset(curve); hb.pp.set_pp(curve, gs, S, H)
    xs = np.stack([O.random_scalars(sf, n, 500 + i) for i in range(b)])
    xs[2] = 0                                   # an all-zero column commits to the identity
    xs[3, 100:] = 0                             # short polynomial (trailing zero coefficients)
    got = hb.group.msm_batch_resident(curve, xs)
    for i in range(b):
        assert (_aff(O, curve, got[i]) == _aff(O, curve, O.msm(curve, gs, xs[i]))).all(), i
    # a sub-range of the generators
    got = hb.group.msm_batch_resident(curve, xs[:, :1000], base_off=512)
    for i in range(b):
        assert (_aff(O, curve, got[i]) == _aff(O, curve, O.msm(curve, gs[512:1512], xs[i, :1000]))).all(), i
    hb.pp.reset(curve)


@pytest.mark.parametrize("kind", ["all_equal", "tiny", "two_values", "top_heavy"])
def test_msm_degenerate_scalar_distributions(hb, oracle, kind):
    """Scalar vectors that put a whole window into a handful of buckets (selector-like columns, constants):
    buckets are cut into bounded tasks, so these must stay exact and must not serialise on one thread."""
    O = oracle
    k = 40000
    pts = O.points_arith(0, k)
    base = O.random_scalars(0, k, 77)
    if kind == "all_equal":
        sc = np.tile(base[0], (k, 1))
    elif kind == "tiny":
        sc = O.ints_to_mont(0, [(i * 7) % 3 for i in range(k)])
    elif kind == "two_values":
        sc = np.where((np.arange(k) % 2 == 0)[:, None], base[1], base[2])
    else:
        p = O.MODULUS[0]
        sc = O.ints_to_mont(0, [p - 1 - (i % 5) for i in range(k)])
    sc = np.ascontiguousarray(sc, dtype=np.uint64)
    got = hb.group.point_dot_affine(0, sc, pts)
    assert (_aff(O, 0, got) == _aff(O, 0, O.msm(0, pts, sc))).all()


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("k", [1, 5, 129, 2048, 5000, 8192])
def test_msm_one_launch_path_and_its_fallback(hb, oracle, curve, k):
    """MSMs up to 2^13 points take the one-launch kernel (msm_small_kernel); a window whose largest bucket is far
    above the mean makes it hand over to the task-balanced pipeline.  Both must give the oracle's point, for uniform
    scalars, for skewed ones (fallback) and with the path switched off (tuning key 1)."""
    from halo_b200._lib import lib
    O = oracle
    sf = O.SCALAR_FIELD[curve]
    pts = O.points_arith(curve, k)
    uniform = O.random_scalars(sf, k, 4000 + k)
    skewed = np.ascontiguousarray(np.tile(uniform[0], (k, 1)), dtype=np.uint64)        # every point in one bucket per window
    half_zero = uniform.copy(); half_zero[::2] = 0
    for sc in (uniform, skewed, half_zero):
        exp = _aff(O, curve, O.msm(curve, pts, sc))
        assert (_aff(O, curve, hb.group.point_dot_affine(curve, sc, pts)) == exp).all()
        lib.halo_set_tuning(1, 1)
        try:
            assert (_aff(O, curve, hb.group.point_dot_affine(curve, sc, pts)) == exp).all()
        finally:
            lib.halo_set_tuning(1, 0)


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("log_k", [21, 22])
def test_msm_beyond_oracle_sizes_known_discrete_logs(hb, oracle, curve, log_k):
    """BASELINE's largest sizes (2^21, 2^22) are out of the CPU oracle's reach in a test; the points are generated as
    m_i*G on the device with known 64-bit m_i, so sum s_i*P_i must equal (sum s_i*m_i mod r)*G — one scalar
    multiplication, checked with the oracle."""
    import ctypes
    O = oracle
    lib = hb._lib.lib
    lib.halo_gen_point_multiplier.restype = ctypes.c_uint64
    sf = O.SCALAR_FIELD[curve]
    r = O.MODULUS[sf]
    k = 1 << log_k
    seed = 77 + log_k
    pts = np.empty((k, 8), dtype=np.uint64)
    hb._lib.check(lib.halo_gen_points(curve, ctypes.c_uint64(seed), ctypes.c_size_t(0), ctypes.c_size_t(k), hb._lib.ptr(pts.reshape(-1))))
    for i in (0, 1, k - 1):                              # spot-check the generator against the oracle
        m = lib.halo_gen_point_multiplier(ctypes.c_uint64(seed), ctypes.c_uint64(i))
        e = O.point_mul(curve, O.generator(curve), O.ints_to_mont(sf, [m])[0])
        assert (_aff(O, curve, e)[0] == pts[i]).all()
    sc = O.random_scalars(sf, k, log_k)
    got = hb.group.point_dot_affine(curve, sc, pts)
    # sum s_i m_i with numpy object arithmetic (scalars are Montgomery limbs: value * R^-1)
    with np.errstate(over="ignore"):
        x = np.uint64(seed) + np.arange(k, dtype=np.uint64) * np.uint64(0x2545F4914F6CDD1D) + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        ms = (x ^ (x >> np.uint64(31))) | np.uint64(1)
    s = sc.astype(object)
    vals = s[:, 0] + (s[:, 1] << 64) + (s[:, 2] << 128) + (s[:, 3] << 192)
    t = int((vals * ms.astype(object)).sum()) % r * pow(1 << 256, -1, r) % r
    exp = O.point_mul(curve, O.generator(curve), O.ints_to_mont(sf, [t])[0])
    assert (_aff(O, curve, got) == _aff(O, curve, exp)).all()


@pytest.mark.parametrize("curve", [0, 1])
@pytest.mark.parametrize("G", [64, 128, 256])
def test_msm_batched_affine_path_forced(hb, oracle, curve, G):
    """The batched-affine bucket accumulation (msm_accumulate_ba, tuning key 8) forced at sizes the oracle handles, on the
    inputs that exercise every case of the affine group law inside a batch: identity bases, duplicates (P + P),
    negated pairs with equal scalars (P - P), all-equal scalars (one bucket per window, spanning many thread
    boundaries: head / tail partial sums), tiny scalars, ragged sizes around the per-thread range.  Results must equal
    the oracle's and the task path's bit for bit."""
    O = oracle
    lib = hb._lib.lib
    sf, bf = O.SCALAR_FIELD[curve], O.BASE_FIELD[curve]

    def run(sc, pts):
        lib.halo_set_tuning(8, G); lib.halo_set_tuning(1, 1)          # batched affine on, one-launch small path off
        try:
            got = hb.group.point_dot_affine(curve, sc, pts)
        finally:
            lib.halo_set_tuning(8, 0); lib.halo_set_tuning(1, 0)
        exp = O.msm(curve, pts, sc)
        assert (_aff(O, curve, got) == _aff(O, curve, exp)).all()
        old = hb.group.point_dot_affine(curve, sc, pts)                # the default task path
        assert (_aff(O, curve, old) == _aff(O, curve, exp)).all()

    for k in (1, 2, 3, G - 1, G, G + 1, 1000, 5000, (1 << 14) + 5):
        run(O.random_scalars(sf, k, 4000 + k + G), O.points_arith(curve, k, k0=0x1000 + k))
    k = 6000
    pts = O.points_arith(curve, k).copy()
    neg_y = lambda y: O.field_sub(bf, np.zeros((1, 4), dtype=np.uint64), y.reshape(1, 4))[0]
    pts[17] = 0; pts[300:340] = 0                                      # identity bases
    pts[1000:1400] = pts[1000]                                         # 400 copies of one point
    pts[2000:2200:2] = pts[2001:2200:2]                                # adjacent duplicates
    for i in range(3000, 3100, 2):                                     # P, -P pairs
        pts[i + 1, 0:4] = pts[i, 0:4]; pts[i + 1, 4:8] = neg_y(pts[i, 4:8])
    sc = O.random_scalars(sf, k, 555 + G)
    sc[1000:1400] = sc[1000]                                           # equal scalars on equal points: doublings in every round
    sc[2000:2200:2] = sc[2001:2200:2]
    sc[3000:3100:2] = sc[3001:3100:2]                                  # s P + s (-P) = 0
    sc[50:60] = 0
    run(sc, pts)
    run(np.ascontiguousarray(np.tile(sc[7], (k, 1))), pts)             # all equal: one bucket per window
    run(O.ints_to_mont(sf, [(i * 7) % 3 for i in range(k)]), pts)      # digits 0, 1, 2 only
    p = O.MODULUS[sf]
    run(O.ints_to_mont(sf, [p - 1 - (i % 5) for i in range(k)]), pts)  # top-heavy
    run(O.edge_scalars(sf, k, 98), pts)
    same = np.ascontiguousarray(np.tile(pts[5], (k, 1)))               # ONE point k times: a pure doubling tree
    run(np.ascontiguousarray(np.tile(sc[9], (k, 1))), same)


@pytest.mark.parametrize("curve", [0, 1])
def test_msm_host_buffers_pipelined_by_point_range(hb, oracle, curve):
    """halo_msm over host buffers of >= 2^19 points uploads and accumulates by point range into one bucket array
    (msm_host_chunked): 4 ranges from pinned memory, 2 staged ranges from pageable memory.  Both must equal the oracle
    and the unchunked path (tuning key 9 = 1), on a ragged size with identity / duplicate / negated bases placed across
    the range boundaries and scalars that leave some ranges' buckets empty."""
    import ctypes
    import torch
    O = oracle
    lib = hb._lib.lib
    sf, bf = O.SCALAR_FIELD[curve], O.BASE_FIELD[curve]
    k = (1 << 19) + 37
    pts = O.points_arith(curve, k, k0=0x515 + curve).copy()
    sc = O.random_scalars(sf, k, 6100 + curve)
    q = k // 4
    pts[q - 1] = pts[q]; sc[q - 1] = sc[q]                                    # the same term on both sides of a boundary
    pts[2 * q] = pts[5]; pts[2 * q, 4:8] = O.field_sub(bf, np.zeros((1, 4), dtype=np.uint64), pts[5, 4:8].reshape(1, 4))[0]
    sc[2 * q] = sc[5]                                                         # s P in range 0, s (-P) in range 2
    pts[3 * q: 3 * q + 50] = 0                                                # identity bases
    sc[q: q + 20000] = sc[7]                                                  # one bucket per window for part of range 1
    sc[k - 5000:] = 0                                                         # nothing to add at the end of the last range
    exp = _aff(O, curve, O.msm(curve, pts, sc))
    got = hb.group.point_dot_affine(curve, sc, pts)                           # pageable numpy buffers: 2 staged ranges
    assert (_aff(O, curve, got) == exp).all()
    hp = torch.from_numpy(pts.view(np.int64)).pin_memory()
    hs = torch.from_numpy(sc.view(np.int64)).pin_memory()
    u64p = ctypes.POINTER(ctypes.c_uint64)
    out = np.zeros(12, dtype=np.uint64)

    def pinned():
        hb._lib.check(lib.halo_msm(curve, ctypes.cast(hp.data_ptr(), u64p), ctypes.c_size_t(0), ctypes.cast(hs.data_ptr(), u64p),
                                   ctypes.c_size_t(k), out.ctypes.data_as(u64p)))
        return out.copy()
    assert (_aff(O, curve, pinned()) == exp).all()                            # 4 ranges, asynchronous copies
    assert (_aff(O, curve, pinned()) == exp).all()                            # again: the bucket array starts from scratch
    lib.halo_set_tuning(9, 1)
    try:
        assert (_aff(O, curve, pinned()) == exp).all()                        # unchunked
        assert (_aff(O, curve, hb.group.point_dot_affine(curve, sc, pts)) == exp).all()
    finally:
        lib.halo_set_tuning(9, 0)


@pytest.mark.parametrize("curve", [0, 1])
def test_commit_over_resident_generators_fixed_base_table(hb, oracle, curve):
    """Commitments over a large resident parameter set take the fixed-base table (rows 2^(c w) * Gs[i], one bucket window
    per MSM): equal to the ordinary pipeline (tuning key 14 = 1) and to the oracle, for the full length, a ragged prefix
    above half the set, a short prefix (ordinary pipeline), a batch, and scalar vectors that stress the digit columns
    (all p - 1: every digit at its signed extreme with carries into the top column; 0 / 1; a duplicate-heavy generator
    set with identities as in the reference's)."""
    O = oracle
    lib = hb._lib.lib
    sf = O.SCALAR_FIELD[curve]
    p = O.MODULUS[sf]
    n = 1 << 18
    gs = O.points_arith(curve, n + 2, k0=0x1234 + curve)
    gs[5] = gs[4]; gs[7] = 0; gs[n - 1] = gs[0]                      # duplicates and an identity among the generators
    hb.pp.reset(curve)
    hb.pp.set_pp(curve, gs[:n], gs[n], gs[n + 1])
    aff = lambda q: O.point_to_affine(curve, np.asarray(q, dtype=np.uint64).reshape(-1, 12))
    try:
        cases = {"random": O.random_scalars(sf, n, 5), "pm1": np.tile(O.ints_to_mont(sf, [p - 1]), (n, 1)),
                 "zero_one": O.ints_to_mont(sf, [0, 1] * (n // 2))}
        cases["random"][3] = 0
        for name, sc in cases.items():
            for k in (n, n // 2 + 7, n // 4):
                got = hb.pcdl.commit(curve, sc[:k], n - 1)
                lib.halo_set_tuning(14, 1)
                try:
                    plain = hb.pcdl.commit(curve, sc[:k], n - 1)
                finally:
                    lib.halo_set_tuning(14, 0)
                assert (aff(got) == aff(plain)).all(), (name, k)
                if name == "random" or k == n:
                    assert (aff(got) == aff(O.msm(curve, gs[:k], sc[:k]))).all(), (name, k)
        # a batch over the resident set (the 16 commits of a prover round use this shape)
        b = 3
        cols = O.random_scalars(sf, b * n, 9).reshape(b, n, 4)
        cols[1, :100] = 0
        got = hb.group.msm_batch_resident(curve, cols) if hasattr(hb.group, "msm_batch_resident") else None
        if got is None:
            import ctypes
            out = np.zeros((b, 12), dtype=np.uint64)
            hb._lib.check(lib.halo_msm_batch(curve, ctypes.c_size_t(0), hb._lib.ptr(cols.reshape(-1)), ctypes.c_size_t(n), ctypes.c_size_t(b),
                                             hb._lib.ptr(out.reshape(-1))))
            got = out
        for j in range(b):
            assert (aff(got[j]) == aff(O.msm(curve, gs[:n], cols[j]))).all(), j
    finally:
        hb.pp.reset(curve)
