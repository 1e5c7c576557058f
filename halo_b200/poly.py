"""crates/group/src/poly.rs — Domain and Evals, with the transforms and the pointwise arithmetic on the GPU.

`Domain(field, n)` stands for `Domain::<P>::new(n)` (= GeneralEvaluationDomain over P's scalar field, poly.rs:11);
`Evals` mirrors the wrapper at poly.rs:14-144 including its index convention: `from_vec_and_domain` rotates the
vector right by one so that user row i sits at omega^(i+1) (poly.rs:21-24).  Polynomials are coefficient arrays
(k,4) with trailing zeros stripped, as ark-poly's DensePolynomial keeps them."""
import ctypes

import numpy as np

from . import _lib, consts, group
from ._lib import lib, check, ptr, as_u64


def _strip(coeffs):
    nz = np.flatnonzero(coeffs.any(axis=1))
    return coeffs[: (nz[-1] + 1 if nz.size else 0)]


class Domain:
    """Radix-2 multiplicative subgroup of size n (next power of two of the request, as ark-poly rounds up)."""

    def __init__(self, field, n):
        if n < 1:
            raise _lib.PreconditionError(_lib.HALO_EINVAL, "domain size must be positive")
        self.field = field
        self.log_n = max(0, (int(n) - 1).bit_length())
        self.n = 1 << self.log_n
        if self.log_n > 30:
            raise _lib.PreconditionError(_lib.HALO_EINVAL, "domain too large")

    def size(self):
        return self.n

    def __eq__(self, other):
        return isinstance(other, Domain) and (self.field, self.n) == (other.field, other.n)

    def element(self, i):
        """omega^i (Montgomery limbs)."""
        return consts.root_power(self.field, self.log_n, i)


def ntt(field, data, inverse=False, coset=None):
    """In-place-style transform of (b, N, 4) or (N, 4) arrays; returns a new array.  halo_ntt in the C ABI."""
    _lib.ensure_init()
    a = np.ascontiguousarray(data, dtype=np.uint64)
    shape = a.shape
    a = a.reshape(-1, shape[-2], 4).copy()
    n = a.shape[1]
    assert n & (n - 1) == 0, "length must be a power of two"
    g = as_u64(coset, 4).reshape(4) if coset is not None else None
    check(lib.halo_ntt(field, ptr(a.reshape(-1, 4)), ctypes.c_uint(n.bit_length() - 1), ctypes.c_size_t(a.shape[0]),
                       ctypes.c_int(1 if inverse else 0), ptr(g)))
    return a.reshape(shape)


class Evals:
    def __init__(self, evals, domain):
        self.evals = as_u64(evals, 4)
        self.dom = domain
        assert self.evals.shape[0] == domain.n

    # ---- constructors (poly.rs:21-24, 56-72) ----
    @classmethod
    def from_vec_and_domain(cls, evals, domain):
        return cls(evals, domain).shift_right()

    @classmethod
    def from_poly_ref(cls, coeffs, domain):
        """evaluate_over_domain_by_ref: zero-pads short coefficient vectors and reduces long ones mod X^N - 1."""
        _lib.ensure_init()
        c = as_u64(coeffs, 4)
        out = np.empty((domain.n, 4), dtype=np.uint64)
        check(lib.halo_poly_evaluate(domain.field, ptr(c) if c.shape[0] else None, ctypes.c_size_t(c.shape[0]),
                                     ctypes.c_uint(domain.log_n), ptr(out)))
        return cls(out, domain)

    from_poly = from_poly_ref

    @classmethod
    def one(cls, domain):
        return cls.from_vec_and_domain(np.tile(consts.one(domain.field), (domain.n, 1)), domain)

    @classmethod
    def zero(cls, domain):
        return cls.from_vec_and_domain(np.zeros((domain.n, 4), dtype=np.uint64), domain)

    # ---- rotations (poly.rs:26-45) ----
    def shift_right(self):
        assert self.evals.shape[0] > 0
        return Evals(np.roll(self.evals, 1, axis=0), self.dom)

    def shift_left(self):
        return Evals(np.roll(self.evals, -1, axis=0), self.dom)

    def shift_left_small_domain(self, small_domain):
        return Evals(np.roll(self.evals, -(self.dom.n // small_domain.n), axis=0), self.dom)

    def index_small_domain(self, i, small_domain):
        return self.evals[i * (self.dom.n // small_domain.n)]

    def extend(self, new_domain_size):
        v = np.zeros((new_domain_size, 4), dtype=np.uint64)
        v[: self.dom.n] = self.evals
        return Evals.from_vec_and_domain(v, Domain(self.dom.field, new_domain_size))

    # ---- transforms (poly.rs:133-139) ----
    def interpolate(self):
        return _strip(ntt(self.dom.field, self.evals, inverse=True))

    interpolate_by_ref = interpolate

    def domain(self):
        return self.dom

    def omega(self):
        return self.dom.element(1)

    def __getitem__(self, i):
        return self.evals[i]

    # ---- pointwise arithmetic (poly.rs:84-127,157-327), on the GPU ----
    def _bin(self, op, other):
        if isinstance(other, Evals):
            assert other.dom == self.dom, "domains must match"
            b = other.evals
        else:
            b = np.tile(as_u64(other, 4).reshape(1, 4), (self.dom.n, 1))
        return Evals(group.field_op(self.dom.field, op, self.evals, b), self.dom)

    def __add__(self, o):
        return self._bin("add", o)

    def __sub__(self, o):
        return self._bin("sub", o)

    def __mul__(self, o):
        return self._bin("mul", o)

    def scale(self, s):
        return self._bin("mul", s)

    scale_ref = scale

    def add_scalar(self, s):
        return self._bin("add", s)

    add_scalar_ref = add_scalar

    def sub_scalar(self, s):
        return self._bin("sub", s)

    sub_scalar_ref = sub_scalar

    def divide_by_vanishing(self, vanishing_domain):
        """poly.rs:74-82: divide evaluation i by Z_H(omega^i) = omega^(i*|H|) - 1; raises when a point of this
        domain lies in H (the reference unwraps a None inverse there)."""
        f = self.dom.field
        w = np.stack([consts.root_power(f, self.dom.log_n, (i * vanishing_domain.n) % self.dom.n)
                      for i in range(self.dom.n)])
        z = group.field_op(f, "sub", w, np.tile(consts.one(f), (self.dom.n, 1)))
        if not z.any(axis=1).all():
            raise _lib.PreconditionError(_lib.HALO_EINVAL, "vanishing polynomial is zero on this domain")
        return Evals(group.field_op(f, "mul", self.evals, group.field_op(f, "inv", z)), self.dom)

    def __eq__(self, o):
        return isinstance(o, Evals) and self.dom == o.dom and bool((self.evals == o.evals).all())


def poly_mul(field, a, b):
    """≙ `&a * &b` on DensePolynomial (FFT-based product inside ark-poly; protocol.rs:132-139,198-199, pcdl.rs:215).
    Returns the coefficients with trailing zeros stripped."""
    _lib.ensure_init()
    a, b = _strip(as_u64(a, 4)), _strip(as_u64(b, 4))
    if a.shape[0] == 0 or b.shape[0] == 0:
        return np.zeros((0, 4), dtype=np.uint64)
    out = np.empty((a.shape[0] + b.shape[0] - 1, 4), dtype=np.uint64)
    check(lib.halo_poly_mul(field, ptr(a), ctypes.c_size_t(a.shape[0]), ptr(b), ctypes.c_size_t(b.shape[0]), ptr(out)))
    return _strip(out)
