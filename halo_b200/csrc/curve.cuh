// Short-Weierstrass group law for Pallas / Vesta (y^2 = x^3 + 5, a = 0) on top of field.cuh.
//
// Replaces, on the device, what the reference gets from ark-ec 0.5.0 short_weierstrass::{Affine,Projective}
// (call sites: crates/group/src/group.rs:48-56, crates/accumulation/src/pedersen.rs:21-23,
// crates/accumulation/src/pcdl.rs:413-429).  Layouts follow the reference's FFI struct WrappedPoint
// (crates/group/src/wrappers.rs:592-597): affine = {x,y} Montgomery, identity encoded as (0,0); projective results
// are Jacobian (X,Y,Z) like ark's Projective (crates/plonk/src/frontend/ivc/mod.rs:40-45).
//
// Bucket accumulators use extended-Jacobian XYZZ coordinates (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2): a mixed
// addition is 8M + 2S and never needs Z itself.  Every formula takes the doubling / inverse / identity branches:
// the reference's generator set is duplicate-heavy, so P+P and P+(-P) do occur inside buckets.
#pragma once
#include "field.cuh"

namespace halo {

// curve 0 = Pallas: base field halo-Fq (1), scalar field halo-Fp (0); curve 1 = Vesta: the other way round
template <int CID> struct CurveTraits;
template <> struct CurveTraits<0> { static constexpr int BASE = 1, SCALAR = 0; };
template <> struct CurveTraits<1> { static constexpr int BASE = 0, SCALAR = 1; };

template <int CID> struct Affine {
    typedef Fe<CurveTraits<CID>::BASE> F;
    F x, y;
};
template <int CID> struct Jac {
    typedef Fe<CurveTraits<CID>::BASE> F;
    F X, Y, Z;
};
template <int CID> struct XYZZ {
    typedef Fe<CurveTraits<CID>::BASE> F;
    F X, Y, ZZ, ZZZ;
};

#define HALO_BF CurveTraits<CID>::BASE

template <int CID> HD bool affine_is_identity(const Affine<CID>& p) {
    return fe_is_zero<HALO_BF>(p.x) && fe_is_zero<HALO_BF>(p.y);
}
template <int CID> HD XYZZ<CID> xyzz_identity() {
    XYZZ<CID> r;
    r.X = fe_zero<HALO_BF>(); r.Y = fe_zero<HALO_BF>(); r.ZZ = fe_zero<HALO_BF>(); r.ZZZ = fe_zero<HALO_BF>();
    return r;
}
template <int CID> HD bool xyzz_is_identity(const XYZZ<CID>& p) { return fe_is_zero<HALO_BF>(p.ZZ); }

template <int CID> HD XYZZ<CID> xyzz_from_affine(const Affine<CID>& a) {
    XYZZ<CID> r;
    if (affine_is_identity<CID>(a)) return xyzz_identity<CID>();
    r.X = a.x; r.Y = a.y; r.ZZ = fe_one<HALO_BF>(); r.ZZZ = fe_one<HALO_BF>();
    return r;
}

// 2 * (affine point), result XYZZ
template <int CID> HD XYZZ<CID> xyzz_dbl_affine(const Affine<CID>& a) {
    constexpr int B = HALO_BF;
    XYZZ<CID> r;
    if (affine_is_identity<CID>(a) || fe_is_zero<B>(a.y)) return xyzz_identity<CID>();
    Fe<B> U = fe_dbl<B>(a.y);
    Fe<B> V = fe_sqr<B>(U);
    Fe<B> W = fe_mul<B>(U, V);
    Fe<B> S = fe_mul<B>(a.x, V);
    Fe<B> xx = fe_sqr<B>(a.x);
    Fe<B> M = fe_add<B>(fe_dbl<B>(xx), xx);
    r.X = fe_sub<B>(fe_sqr<B>(M), fe_dbl<B>(S));
    r.Y = fe_sub<B>(fe_mul<B>(M, fe_sub<B>(S, r.X)), fe_mul<B>(W, a.y));
    r.ZZ = V; r.ZZZ = W;
    return r;
}

template <int CID> HD XYZZ<CID> xyzz_dbl(const XYZZ<CID>& p) {
    constexpr int B = HALO_BF;
    if (xyzz_is_identity<CID>(p) || fe_is_zero<B>(p.Y)) return xyzz_identity<CID>();
    XYZZ<CID> r;
    Fe<B> U = fe_dbl<B>(p.Y);
    Fe<B> V = fe_sqr<B>(U);
    Fe<B> W = fe_mul<B>(U, V);
    Fe<B> S = fe_mul<B>(p.X, V);
    Fe<B> xx = fe_sqr<B>(p.X);
    Fe<B> M = fe_add<B>(fe_dbl<B>(xx), xx);
    r.X = fe_sub<B>(fe_sqr<B>(M), fe_dbl<B>(S));
    r.Y = fe_sub<B>(fe_mul<B>(M, fe_sub<B>(S, r.X)), fe_mul<B>(W, p.Y));
    r.ZZ = fe_mul<B>(V, p.ZZ);
    r.ZZZ = fe_mul<B>(W, p.ZZZ);
    return r;
}

// The doubling cases of the addition formulas (P + P) are rare; as out-of-line calls they stay out of the hot loops'
// instruction stream (an inlined doubling is ~40 KB of code at every addition site).
#if defined(__CUDACC__)
#define HALO_COLD __host__ __device__ __noinline__
#else
#define HALO_COLD __attribute__((noinline))
#endif
template <int CID> HALO_COLD XYZZ<CID> xyzz_dbl_affine_cold(Affine<CID> a) { return xyzz_dbl_affine<CID>(a); }
template <int CID> HALO_COLD XYZZ<CID> xyzz_dbl_cold(XYZZ<CID> p) { return xyzz_dbl<CID>(p); }

// acc += (negate ? -a : a)   (mixed addition, 8M + 2S on the common path)
template <int CID> HD void xyzz_madd(XYZZ<CID>& acc, const Affine<CID>& a_in, bool negate) {
    constexpr int B = HALO_BF;
    if (affine_is_identity<CID>(a_in)) return;
    Affine<CID> a = a_in;
    if (negate) a.y = fe_neg<B>(a.y);
    if (xyzz_is_identity<CID>(acc)) { acc = xyzz_from_affine<CID>(a); return; }
    Fe<B> U2 = fe_mul<B>(a.x, acc.ZZ);
    Fe<B> S2 = fe_mul<B>(a.y, acc.ZZZ);
    Fe<B> P = fe_sub<B>(U2, acc.X);
    Fe<B> R = fe_sub<B>(S2, acc.Y);
    if (fe_is_zero<B>(P)) {
        if (fe_is_zero<B>(R)) acc = xyzz_dbl_affine_cold<CID>(a);
        else acc = xyzz_identity<CID>();
        return;
    }
    Fe<B> PP = fe_sqr<B>(P);
    Fe<B> PPP = fe_mul<B>(P, PP);
    Fe<B> Q = fe_mul<B>(acc.X, PP);
    Fe<B> X3 = fe_sub<B>(fe_sub<B>(fe_sqr<B>(R), PPP), fe_dbl<B>(Q));
    Fe<B> Y3 = fe_sub<B>(fe_mul<B>(R, fe_sub<B>(Q, X3)), fe_mul<B>(acc.Y, PPP));
    acc.X = X3; acc.Y = Y3;
    acc.ZZ = fe_mul<B>(acc.ZZ, PP);
    acc.ZZZ = fe_mul<B>(acc.ZZZ, PPP);
}

// acc += o   (general addition, 12M + 2S)
template <int CID> HD void xyzz_add(XYZZ<CID>& acc, const XYZZ<CID>& o) {
    constexpr int B = HALO_BF;
    if (xyzz_is_identity<CID>(o)) return;
    if (xyzz_is_identity<CID>(acc)) { acc = o; return; }
    Fe<B> U1 = fe_mul<B>(acc.X, o.ZZ);
    Fe<B> U2 = fe_mul<B>(o.X, acc.ZZ);
    Fe<B> S1 = fe_mul<B>(acc.Y, o.ZZZ);
    Fe<B> S2 = fe_mul<B>(o.Y, acc.ZZZ);
    Fe<B> P = fe_sub<B>(U2, U1);
    Fe<B> R = fe_sub<B>(S2, S1);
    if (fe_is_zero<B>(P)) {
        if (fe_is_zero<B>(R)) acc = xyzz_dbl_cold<CID>(acc);
        else acc = xyzz_identity<CID>();
        return;
    }
    Fe<B> PP = fe_sqr<B>(P);
    Fe<B> PPP = fe_mul<B>(P, PP);
    Fe<B> Q = fe_mul<B>(U1, PP);
    Fe<B> X3 = fe_sub<B>(fe_sub<B>(fe_sqr<B>(R), PPP), fe_dbl<B>(Q));
    Fe<B> Y3 = fe_sub<B>(fe_mul<B>(R, fe_sub<B>(Q, X3)), fe_mul<B>(S1, PPP));
    acc.X = X3; acc.Y = Y3;
    acc.ZZ = fe_mul<B>(fe_mul<B>(acc.ZZ, o.ZZ), PP);
    acc.ZZZ = fe_mul<B>(fe_mul<B>(acc.ZZZ, o.ZZZ), PPP);
}

// XYZZ -> Jacobian without inversion: (X*ZZ^2, Y*ZZ^3, ZZZ)
template <int CID> HD Jac<CID> xyzz_to_jac(const XYZZ<CID>& p) {
    constexpr int B = HALO_BF;
    Jac<CID> r;
    if (xyzz_is_identity<CID>(p)) { r.X = fe_one<B>(); r.Y = fe_one<B>(); r.Z = fe_zero<B>(); return r; }
    Fe<B> z2 = fe_sqr<B>(p.ZZ);
    r.X = fe_mul<B>(p.X, z2);
    r.Y = fe_mul<B>(p.Y, fe_mul<B>(z2, p.ZZ));
    r.Z = p.ZZZ;
    return r;
}
// Jacobian -> XYZZ: (X, Y, Z^2, Z^3)
template <int CID> HD XYZZ<CID> jac_to_xyzz(const Jac<CID>& p) {
    constexpr int B = HALO_BF;
    XYZZ<CID> r;
    if (fe_is_zero<B>(p.Z)) return xyzz_identity<CID>();
    r.X = p.X; r.Y = p.Y;
    r.ZZ = fe_sqr<B>(p.Z);
    r.ZZZ = fe_mul<B>(r.ZZ, p.Z);
    return r;
}
// XYZZ -> affine (one inversion); identity -> (0,0)
template <int CID> HD Affine<CID> xyzz_to_affine(const XYZZ<CID>& p) {
    constexpr int B = HALO_BF;
    Affine<CID> a;
    if (xyzz_is_identity<CID>(p)) { a.x = fe_zero<B>(); a.y = fe_zero<B>(); return a; }
    // 1/ZZZ and 1/ZZ from one inversion of ZZ*ZZZ
    Fe<B> t = fe_inv<B>(fe_mul<B>(p.ZZ, p.ZZZ));
    a.x = fe_mul<B>(p.X, fe_mul<B>(t, p.ZZZ));
    a.y = fe_mul<B>(p.Y, fe_mul<B>(t, p.ZZ));
    return a;
}
template <int CID> HD Affine<CID> jac_to_affine(const Jac<CID>& p) { return xyzz_to_affine<CID>(jac_to_xyzz<CID>(p)); }

// k * P for a scalar given as canonical bits (8 x u32), plain double-and-add; host-side O(lg n) work and tests
template <int CID> HD XYZZ<CID> xyzz_mul_bits(const XYZZ<CID>& p, const u32* k, int nbits = 256) {
    XYZZ<CID> acc = xyzz_identity<CID>();
    for (int i = nbits - 1; i >= 0; --i) {
        acc = xyzz_dbl<CID>(acc);
        if ((k[i >> 5] >> (i & 31)) & 1) xyzz_add<CID>(acc, p);
    }
    return acc;
}

// group equality of two XYZZ points
template <int CID> HD bool xyzz_eq(const XYZZ<CID>& a, const XYZZ<CID>& b) {
    constexpr int B = HALO_BF;
    bool ia = xyzz_is_identity<CID>(a), ib = xyzz_is_identity<CID>(b);
    if (ia || ib) return ia && ib;
    return fe_eq<B>(fe_mul<B>(a.X, b.ZZ), fe_mul<B>(b.X, a.ZZ)) && fe_eq<B>(fe_mul<B>(a.Y, b.ZZZ), fe_mul<B>(b.Y, a.ZZZ));
}

}  // namespace halo
