// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked, imported or executed by the
// product path (halo_b200/); only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may use anything in oracle/.
//
// CPU restatement of the prime-field arithmetic that the reference obtains
// from ark-ff 0.5.0 (`Fp<MontBackend<_,4>,4>`; pinned in
// /root/reference/crates/Cargo.lock:115-245, source NOT vendored in the tree).
// Conventions restated:
//   * elements are 4 x u64 little-endian limbs in Montgomery form, R = 2^256
//     (reference layout: crates/group/src/wrappers.rs:603-616, `.0.0` /
//     `new_unchecked(BigInt(..))`)
//   * halo `Fp` = ark_pallas::Fr, halo `Fq` = ark_pallas::Fq
//     (crates/group/src/lib.rs:8-9)
//   * multiplicative generator 5, two-adicity 32 (roots of unity pinned by
//     crates/plonk/src/frontend/ivc/mod.rs:56,113)
// Parity pinning: see oracle/README.md (golden accumulators G1/G2, roots G3,
// Kimchi Poseidon vectors G4).
#pragma once
#include <cstdint>
#include <cstring>
#include <array>

namespace orc {

typedef uint64_t u64;
typedef unsigned __int128 u128;

struct FieldParams {
    u64 p[4];      // modulus
    u64 inv;       // -p^{-1} mod 2^64
    u64 r[4];      // R mod p   (Montgomery one)
    u64 r2[4];     // R^2 mod p
};

// field id 0: halo Fp (Pallas scalar / Vesta base); 1: halo Fq (Pallas base / Vesta scalar)
extern const FieldParams FIELD_PARAMS[2];
static inline const FieldParams& field_params(int id) { return FIELD_PARAMS[id]; }

static inline bool geq4(const u64 a[4], const u64 b[4]) {
    for (int i = 3; i >= 0; --i) {
        if (a[i] != b[i]) return a[i] > b[i];
    }
    return true;
}
static inline u64 add4(u64 out[4], const u64 a[4], const u64 b[4]) {
    u128 c = 0;
    for (int i = 0; i < 4; ++i) { c += (u128)a[i] + b[i]; out[i] = (u64)c; c >>= 64; }
    return (u64)c;
}
static inline u64 sub4(u64 out[4], const u64 a[4], const u64 b[4]) {
    u64 borrow = 0;
    for (int i = 0; i < 4; ++i) {
        u128 d = (u128)a[i] - b[i] - borrow;
        out[i] = (u64)d; borrow = (u64)(d >> 64) & 1;
    }
    return borrow;
}

template <int FID>
struct Fe {
    u64 l[4];

    static const FieldParams& P() { return field_params(FID); }

    static Fe zero() { Fe r; memset(r.l, 0, 32); return r; }
    static Fe one() { Fe r; memcpy(r.l, P().r, 32); return r; }
    static Fe from_raw(const u64* w) { Fe r; memcpy(r.l, w, 32); return r; }  // already Montgomery
    void to_raw(u64* w) const { memcpy(w, l, 32); }

    bool is_zero() const { return (l[0] | l[1] | l[2] | l[3]) == 0; }
    bool operator==(const Fe& o) const { return memcmp(l, o.l, 32) == 0; }
    bool operator!=(const Fe& o) const { return !(*this == o); }

    Fe operator+(const Fe& o) const {
        Fe r; u64 c = add4(r.l, l, o.l);
        if (c || geq4(r.l, P().p)) sub4(r.l, r.l, P().p);
        return r;
    }
    Fe operator-(const Fe& o) const {
        Fe r; u64 b = sub4(r.l, l, o.l);
        if (b) add4(r.l, r.l, P().p);
        return r;
    }
    Fe neg() const { return zero() - *this; }
    Fe dbl() const { return *this + *this; }

    // CIOS Montgomery multiplication, 64-bit limbs.
    Fe operator*(const Fe& o) const {
        const FieldParams& F = P();
        u64 t[6] = {0, 0, 0, 0, 0, 0};
        for (int i = 0; i < 4; ++i) {
            u128 c = 0;
            for (int j = 0; j < 4; ++j) {
                c += (u128)l[j] * o.l[i] + t[j];
                t[j] = (u64)c; c >>= 64;
            }
            c += t[4]; t[4] = (u64)c; t[5] = (u64)(c >> 64);
            u64 m = t[0] * F.inv;
            c = (u128)m * F.p[0] + t[0]; c >>= 64;
            for (int j = 1; j < 4; ++j) {
                c += (u128)m * F.p[j] + t[j];
                t[j - 1] = (u64)c; c >>= 64;
            }
            c += t[4]; t[3] = (u64)c; c >>= 64;
            t[4] = t[5] + (u64)c;
        }
        Fe r; memcpy(r.l, t, 32);
        if (t[4] || geq4(r.l, F.p)) sub4(r.l, r.l, F.p);
        return r;
    }
    Fe sqr() const { return *this * *this; }
    Fe& operator+=(const Fe& o) { *this = *this + o; return *this; }
    Fe& operator-=(const Fe& o) { *this = *this - o; return *this; }
    Fe& operator*=(const Fe& o) { *this = *this * o; return *this; }

    // canonical (non-Montgomery) integer <-> element
    static Fe from_canonical(const u64 w[4]) {   // w < p required
        Fe a; memcpy(a.l, w, 32);
        Fe r2; memcpy(r2.l, P().r2, 32);
        return a * r2;
    }
    void to_canonical(u64 w[4]) const {
        Fe o; o.l[0] = 1; o.l[1] = o.l[2] = o.l[3] = 0;
        Fe r = *this * o;
        memcpy(w, r.l, 32);
    }
    static Fe from_u64(u64 x) { u64 w[4] = {x, 0, 0, 0}; return from_canonical(w); }

    Fe pow(const u64 e[4]) const {
        Fe acc = one();
        for (int i = 255; i >= 0; --i) {
            acc = acc.sqr();
            if ((e[i / 64] >> (i % 64)) & 1) acc = acc * *this;
        }
        return acc;
    }
    Fe pow_u64(u64 e) const { u64 w[4] = {e, 0, 0, 0}; return pow(w); }
    // Fermat inverse; inverse of zero is zero (callers that must panic check first).
    Fe inv() const {
        u64 e[4]; u64 two[4] = {2, 0, 0, 0};
        sub4(e, P().p, two);
        return pow(e);
    }
};

typedef Fe<0> Fp;   // halo naming: Pallas scalar field
typedef Fe<1> Fq;   // halo naming: Pallas base field

// 2^32-th primitive root of unity: 5^((p-1)/2^32); omega_N = that ^ (2^(32-logN)).
template <int FID>
Fe<FID> root_of_unity(unsigned log_n) {
    typedef Fe<FID> F;
    u64 e[4]; u64 one[4] = {1, 0, 0, 0};
    sub4(e, F::P().p, one);
    // e >>= 32
    for (int i = 0; i < 4; ++i) {
        u64 lo = e[i] >> 32;
        u64 hi = (i < 3) ? (e[i + 1] << 32) : 0;
        e[i] = lo | hi;
    }
    F w = F::from_u64(5).pow(e);
    for (unsigned i = log_n; i < 32; ++i) w = w.sqr();
    return w;
}

// batch inversion (Montgomery trick); zeros stay zero.
template <class F>
void batch_inverse(F* v, size_t n) {
    if (n == 0) return;
    F* pre = new F[n];
    F acc = F::one();
    for (size_t i = 0; i < n; ++i) {
        pre[i] = acc;
        if (!v[i].is_zero()) acc = acc * v[i];
    }
    F inv = acc.inv();
    for (size_t i = n; i-- > 0;) {
        if (v[i].is_zero()) continue;
        F t = inv * pre[i];
        inv = inv * v[i];
        v[i] = t;
    }
    delete[] pre;
}

}  // namespace orc
