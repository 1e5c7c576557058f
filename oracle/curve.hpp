// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// CPU restatement of the short-Weierstrass group law the reference obtains
// from ark-ec 0.5.0 (`short_weierstrass::{Affine,Projective}`; Cargo.lock pin,
// source not vendored).  Curves: y^2 = x^3 + 5 over the Pasta cycle
// (SURVEY.md §0).  curve 0 = Pallas (base field halo-Fq, scalars halo-Fp),
// curve 1 = Vesta (base field halo-Fp, scalars halo-Fq) — see
// crates/group/src/wrappers.rs:468-588.
// Points are compared as group elements / canonical affine coordinates, as the
// reference does (crates/poseidon/src/outer_sponge.rs:45-48 hashes after
// into_affine), so any complete formula set is a sound oracle.
#pragma once
#include "field.hpp"
#include <vector>

namespace orc {

template <int CID> struct CurveTraits;
template <> struct CurveTraits<0> { enum { BASE = 1, SCALAR = 0 }; };
template <> struct CurveTraits<1> { enum { BASE = 0, SCALAR = 1 }; };

template <int CID>
struct Affine {
    typedef Fe<CurveTraits<CID>::BASE> F;
    F x, y;
    bool inf;
    static Affine identity() { Affine a; a.x = F::zero(); a.y = F::zero(); a.inf = true; return a; }
    bool on_curve() const {
        if (inf) return true;
        return y.sqr() == x.sqr() * x + F::from_u64(5);
    }
    Affine neg() const { Affine a = *this; if (!inf) a.y = y.neg(); return a; }
    // FFI layout (WrappedPoint, wrappers.rs:592-597): identity encoded as (0,0)
    static Affine from_raw(const u64* xy) {
        Affine a; a.x = F::from_raw(xy); a.y = F::from_raw(xy + 4);
        a.inf = a.x.is_zero() && a.y.is_zero();
        return a;
    }
    void to_raw(u64* xy) const {
        if (inf) { memset(xy, 0, 64); return; }
        x.to_raw(xy); y.to_raw(xy + 4);
    }
};

template <int CID>
struct Jac {
    typedef Fe<CurveTraits<CID>::BASE> F;
    typedef Fe<CurveTraits<CID>::SCALAR> S;
    F X, Y, Z;

    static Jac identity() { Jac p; p.X = F::one(); p.Y = F::one(); p.Z = F::zero(); return p; }
    static Jac from_affine(const Affine<CID>& a) {
        if (a.inf) return identity();
        Jac p; p.X = a.x; p.Y = a.y; p.Z = F::one(); return p;
    }
    static Jac generator() {
        Affine<CID> g; g.x = F::one().neg(); g.y = F::from_u64(2); g.inf = false;
        return from_affine(g);
    }
    static Jac from_raw(const u64* xyz) {
        Jac p; p.X = F::from_raw(xyz); p.Y = F::from_raw(xyz + 4); p.Z = F::from_raw(xyz + 8); return p;
    }
    void to_raw(u64* xyz) const { X.to_raw(xyz); Y.to_raw(xyz + 4); Z.to_raw(xyz + 8); }
    bool is_identity() const { return Z.is_zero(); }
    Jac neg() const { Jac p = *this; p.Y = Y.neg(); return p; }

    Jac dbl() const {
        if (is_identity()) return *this;
        // a = 0 doubling
        F A = X.sqr();
        F B = Y.sqr();
        F C = B.sqr();
        F D = ((X + B).sqr() - A - C).dbl();
        F E = A + A + A;
        F Fv = E.sqr();
        Jac r;
        r.X = Fv - D.dbl();
        r.Y = E * (D - r.X) - C.dbl().dbl().dbl();
        r.Z = (Y * Z).dbl();
        return r;
    }

    Jac add(const Jac& o) const {
        if (is_identity()) return o;
        if (o.is_identity()) return *this;
        F Z1Z1 = Z.sqr(), Z2Z2 = o.Z.sqr();
        F U1 = X * Z2Z2, U2 = o.X * Z1Z1;
        F S1 = Y * o.Z * Z2Z2, S2 = o.Y * Z * Z1Z1;
        if (U1 == U2) {
            if (S1 == S2) return dbl();
            return identity();
        }
        F H = U2 - U1;
        F R = S2 - S1;
        F HH = H.sqr();
        F HHH = H * HH;
        F V = U1 * HH;
        Jac r;
        r.X = R.sqr() - HHH - V.dbl();
        r.Y = R * (V - r.X) - S1 * HHH;
        r.Z = Z * o.Z * H;
        return r;
    }

    Jac add_affine(const Affine<CID>& a) const {
        if (a.inf) return *this;
        if (is_identity()) return from_affine(a);
        F Z1Z1 = Z.sqr();
        F U2 = a.x * Z1Z1;
        F S2 = a.y * Z * Z1Z1;
        if (X == U2) {
            if (Y == S2) return dbl();
            return identity();
        }
        F H = U2 - X;
        F R = S2 - Y;
        F HH = H.sqr();
        F HHH = H * HH;
        F V = X * HH;
        Jac r;
        r.X = R.sqr() - HHH - V.dbl();
        r.Y = R * (V - r.X) - Y * HHH;
        r.Z = Z * H;
        return r;
    }

    // scalar given as canonical 256-bit integer
    Jac mul_bits(const u64 k[4]) const {
        Jac acc = identity();
        for (int i = 255; i >= 0; --i) {
            acc = acc.dbl();
            if ((k[i / 64] >> (i % 64)) & 1) acc = acc.add(*this);
        }
        return acc;
    }
    Jac mul(const S& s) const { u64 k[4]; s.to_canonical(k); return mul_bits(k); }

    Affine<CID> to_affine() const {
        if (is_identity()) return Affine<CID>::identity();
        F zi = Z.inv();
        F zi2 = zi.sqr();
        Affine<CID> a; a.x = X * zi2; a.y = Y * zi2 * zi; a.inf = false;
        return a;
    }
    bool eq(const Jac& o) const {
        if (is_identity() || o.is_identity()) return is_identity() && o.is_identity();
        F Z1Z1 = Z.sqr(), Z2Z2 = o.Z.sqr();
        if (X * Z2Z2 != o.X * Z1Z1) return false;
        return Y * o.Z * Z2Z2 == o.Y * Z * Z1Z1;
    }
};

// CurveGroup::normalize_batch (group.rs:54)
template <int CID>
std::vector<Affine<CID>> normalize_batch(const std::vector<Jac<CID>>& ps) {
    typedef typename Jac<CID>::F F;
    std::vector<F> zs(ps.size());
    for (size_t i = 0; i < ps.size(); ++i) zs[i] = ps[i].Z;
    batch_inverse(zs.data(), zs.size());
    std::vector<Affine<CID>> out(ps.size());
    for (size_t i = 0; i < ps.size(); ++i) {
        if (ps[i].is_identity()) { out[i] = Affine<CID>::identity(); continue; }
        F zi2 = zs[i].sqr();
        out[i].x = ps[i].X * zi2;
        out[i].y = ps[i].Y * zi2 * zs[i];
        out[i].inf = false;
    }
    return out;
}

}  // namespace orc
