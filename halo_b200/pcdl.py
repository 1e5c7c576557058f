"""crates/accumulation/src/pcdl.rs — PC_DL: commit / open / succinct_check / check, Instance, EvalProof, HPoly.

Host mirror over the C ABI.  An `Instance` wraps the flat word layout of include/halo_b200.h
(C[12] d[1] z[4] v[4] | Ls Rs U c hiding C_bar w_prime); field accessors slice it."""
import ctypes

import numpy as np

from . import _lib, consts, pp as _pp
from ._lib import lib, check, ptr, as_u64

lib.halo_instance_words.restype = ctypes.c_size_t
lib.halo_proof_words.restype = ctypes.c_size_t


def _lg(n):
    return n.bit_length() - 1


def setup(curve, gs_xy, S_xy, H_xy):
    """≙ pcdl::setup(n) (pcdl.rs:266-268) with the decoded parameters supplied by the caller."""
    return _pp.set_pp(curve, gs_xy, S_xy, H_xy)


def commit(curve, coeffs, d, w=None):
    """≙ pcdl::commit (pcdl.rs:275-287).  Panics of the reference -> PreconditionError."""
    _lib.ensure_init()
    c = as_u64(coeffs, 4)
    out = np.empty(12, dtype=np.uint64)
    wv = as_u64(w, 4).reshape(4) if w is not None else None
    check(lib.halo_pcdl_commit(curve, ptr(c) if c.shape[0] else None, ctypes.c_size_t(c.shape[0]), ctypes.c_size_t(d),
                               ptr(wv), ptr(out)))
    return out


def chunked_commit(curve, coeffs, d, w=None, chunk_size=1):
    """≙ pcdl::chunked_commit (pcdl.rs:294-314): one commit per `chunk_size` coefficients; returns (count, 12)."""
    _lib.ensure_init()
    c = as_u64(coeffs, 4)
    if chunk_size <= 0:
        raise _lib.PreconditionError(_lib.HALO_EINVAL, "chunk size must be non-zero (slice::chunks panics)")
    cap = -(-c.shape[0] // chunk_size)
    out = np.zeros((max(cap, 1), 12), dtype=np.uint64)
    cnt = ctypes.c_size_t(0)
    wv = as_u64(w, 4).reshape(4) if w is not None else None
    check(lib.halo_pcdl_chunked_commit(curve, ptr(c) if c.shape[0] else None, ctypes.c_size_t(c.shape[0]), ctypes.c_size_t(d),
                                       ptr(wv), ctypes.c_size_t(chunk_size), ptr(out), ctypes.byref(cnt)))
    return out[:cnt.value]


def fold_table_build(curve, n):
    """Explicit set-up of the fixed-base table of the first generator fold for openings of size n (API addition)."""
    _lib.ensure_init()
    check(lib.halo_ipa_fold_table_build(curve, ctypes.c_size_t(n)))


class HPoly:
    """h(X) = prod_{i<lg n} (1 + xi_{lg n - i} X^{2^i}) kept as its challenges (pcdl.rs:191-233)."""

    def __init__(self, field, xis):
        self.field, self.xis = field, as_u64(xis, 4)

    @property
    def lg_n(self):
        return self.xis.shape[0] - 1

    def eval(self, z):
        out = np.empty(4, dtype=np.uint64)
        check(lib.halo_hpoly_eval(self.field, ptr(self.xis), ctypes.c_size_t(self.lg_n), ptr(as_u64(z, 4).reshape(4)), ptr(out)))
        return out

    def get_poly(self):
        _lib.ensure_init()
        out = np.empty((1 << self.lg_n, 4), dtype=np.uint64)
        one = consts.one(self.field).reshape(1, 4)
        check(lib.halo_acc_hpoly(self.field, ptr(self.xis), ptr(one), ctypes.c_size_t(1), ctypes.c_size_t(self.lg_n), ptr(out)))
        return out


class EvalProof:
    def __init__(self, words, lg_n):
        self.w, self.lg_n = np.ascontiguousarray(words, dtype=np.uint64), lg_n

    Ls = property(lambda s: s.w[: 12 * s.lg_n].reshape(-1, 12))
    Rs = property(lambda s: s.w[12 * s.lg_n: 24 * s.lg_n].reshape(-1, 12))
    U = property(lambda s: s.w[24 * s.lg_n: 24 * s.lg_n + 12])
    c = property(lambda s: s.w[24 * s.lg_n + 12: 24 * s.lg_n + 16])
    hiding = property(lambda s: bool(s.w[24 * s.lg_n + 16]))
    C_bar = property(lambda s: s.w[24 * s.lg_n + 17: 24 * s.lg_n + 29] if s.hiding else None)
    w_prime = property(lambda s: s.w[24 * s.lg_n + 29: 24 * s.lg_n + 33] if s.hiding else None)


def open_without_eval(curve, coeffs, C, d, z, v, w=None, hide_q=None, hide_wbar=None, _entry=None):
    """≙ pcdl::open_without_eval (pcdl.rs:326-453); v=None gives pcdl::open (pcdl.rs:463-473).
    Returns (EvalProof, v).  Hiding needs the prover's randomness (the reference draws it from `rng`)."""
    _lib.ensure_init()
    entry = _entry or lib.halo_pcdl_open
    c = as_u64(coeffs, 4)
    n = d + 1
    lg_n = max(_lg(n), 0)
    proof = np.zeros(int(lib.halo_proof_words(lg_n)), dtype=np.uint64)
    vout = np.empty(4, dtype=np.uint64)
    a4 = lambda x: as_u64(x, 4).reshape(4) if x is not None else None
    q = as_u64(hide_q, 4) if hide_q is not None else None
    check(entry(curve, ptr(c) if c.shape[0] else None, ctypes.c_size_t(c.shape[0]),
                ptr(as_u64(C, 12).reshape(12)), ctypes.c_size_t(d), ptr(a4(z)), ptr(a4(v)), ptr(a4(w)),
                ptr(q), ptr(a4(hide_wbar)), ptr(proof), ptr(vout)))
    return EvalProof(proof, lg_n), vout


def open(curve, coeffs, C, d, z, w=None, hide_q=None, hide_wbar=None):
    return open_without_eval(curve, coeffs, C, d, z, None, w, hide_q, hide_wbar)[0]


class Instance:
    """≙ pcdl::Instance {C, d, z, v, pi} (pcdl.rs:28-120)."""

    def __init__(self, curve, words):
        self.curve = curve
        self.w = np.ascontiguousarray(words, dtype=np.uint64)
        self.d = int(self.w[12])
        self.lg_n = _lg(self.d + 1)
        assert self.w.shape[0] == int(lib.halo_instance_words(self.lg_n))

    @classmethod
    def new(cls, curve, C, d, z, v, pi):
        return cls(curve, np.concatenate([as_u64(C, 12).reshape(12), np.array([d], dtype=np.uint64),
                                          as_u64(z, 4).reshape(4), as_u64(v, 4).reshape(4), pi.w]))

    @classmethod
    def open(cls, curve, coeffs, d, z, w=None, hide_q=None, hide_wbar=None):
        """≙ Instance::open (pcdl.rs:41-52): C = commit(p, d, w); v = p(z); pi = open(...)."""
        C = commit(curve, coeffs, d, w)
        pi, v = open_without_eval(curve, coeffs, C, d, z, None, w, hide_q, hide_wbar)
        return cls.new(curve, C, d, z, v, pi)

    @classmethod
    def zero(cls, curve, n):
        """≙ Instance::zero (pcdl.rs:54-65): the zero polynomial opened at z = 0 without hiding."""
        return cls.open(curve, np.zeros((0, 4), dtype=np.uint64), n - 1, np.zeros(4, dtype=np.uint64))

    C = property(lambda s: s.w[0:12])
    z = property(lambda s: s.w[13:17])
    v = property(lambda s: s.w[17:21])
    pi = property(lambda s: EvalProof(s.w[21:], s.lg_n))

    def succinct_check(self):
        """≙ pcdl::succinct_check (pcdl.rs:483-554): returns (HPoly, U) or raises Rejected."""
        _lib.ensure_init()
        xis = np.empty((self.lg_n + 1, 4), dtype=np.uint64)
        U = np.empty(12, dtype=np.uint64)
        check(lib.halo_pcdl_succinct_check(self.curve, ptr(self.w), ptr(xis), ptr(U)))
        return HPoly(consts.SCALAR_FIELD[self.curve], xis), U

    def check(self):
        """≙ pcdl::check (pcdl.rs:563-583): raises Rejected unless the proof is valid."""
        _lib.ensure_init()
        check(lib.halo_pcdl_check(self.curve, ptr(self.w)))
