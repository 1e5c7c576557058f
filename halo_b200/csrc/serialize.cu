// ark-serialize 0.5 "compressed" wire formats of the PC_DL / AS_DL objects, on the host.
//
// Stands in for the derived CanonicalSerialize / CanonicalDeserialize of
//   crates/accumulation/src/pcdl.rs:26-34   Instance  { C, d, z, v, pi }
//   crates/accumulation/src/pcdl.rs:159-168 EvalProof { Ls, Rs, U, c, C_bar, w_prime }
//   crates/accumulation/src/acc.rs:22-26    Accumulator { q: Instance }
// as written by crates/accumulation/src/main.rs:59-91 (qs.bin = Vec<(usize, Instance, Accumulator)>).
// Encodings (ark-serialize / ark-ff / ark-ec 0.5.0, un-vendored; conventions pinned by tests/test_serialize.py on the
// reference's own qs.bin bytes):
//   scalar      32 bytes, little-endian canonical integer (not Montgomery); must be < modulus
//   point       Projective is written as its affine form, compressed: 32 bytes LE canonical x, then one flag byte:
//               bit 7 = y is the larger of the two roots (y > -y as integers), bit 6 = point at infinity (x = 0);
//               both bits set is invalid; bits 0..5 of that byte are ignored on input, as ark-ff does
//   usize       u64 LE          Vec<T>  u64 LE length, then the elements          Option<T>  one byte 0 / 1, then T
// Decompression needs a square root in a field of 2-adicity 32: Tonelli-Shanks with the 2^32-th root of unity 5^t.
// O(lg n) points per object: host code, no device work.
#include "runtime.cuh"
#include "host.cuh"
#include <cstring>

namespace halo {

template <int F> static Fe<F> two_adic_root() {               // 5^((p-1)/2^32): generator of the 2^32-th roots of unity
    u32 e[8];
    for (int i = 0; i < 7; ++i) e[i] = modulus_limb<F>(i + 1);
    e[7] = 0;
    return fe_pow<F>(fe_from_u64<F>(5), e, 256);
}
// a square root of a, or false when a is not a square
template <int F> static bool fe_sqrt(const Fe<F>& a, Fe<F>& out) {
    if (fe_is_zero<F>(a)) { out = a; return true; }
    const Fe<F> one = fe_one<F>();
    u32 e[8];                                                 // (t - 1) / 2 = (p - 1) >> 33 for t = (p - 1) >> 32 odd
    for (int i = 0; i < 8; ++i) {
        const u64 lo = i + 1 < 8 ? modulus_limb<F>(i + 1) : 0u, hi = i + 2 < 8 ? modulus_limb<F>(i + 2) : 0u;
        e[i] = (u32)(((hi << 32) | lo) >> 1);
    }
    Fe<F> w = fe_pow<F>(a, e, 256);
    Fe<F> x = fe_mul<F>(a, w), b = fe_mul<F>(x, w), z = two_adic_root<F>();
    unsigned v = 32;
    while (!fe_eq<F>(b, one)) {
        unsigned k = 0;
        Fe<F> b2 = b;
        while (!fe_eq<F>(b2, one)) { b2 = fe_sqr<F>(b2); if (++k == v) return false; }
        w = z;
        for (unsigned i = 0; i + k + 1 < v; ++i) w = fe_sqr<F>(w);
        z = fe_sqr<F>(w);
        b = fe_mul<F>(b, z);
        x = fe_mul<F>(x, w);
        v = k;
    }
    if (!fe_eq<F>(fe_sqr<F>(x), a)) return false;
    out = x;
    return true;
}
template <int F> static bool canon_less(const Fe<F>& a, const Fe<F>& b) {   // canonical integers, a < b
    for (int i = 7; i >= 0; --i) if (a.l[i] != b.l[i]) return a.l[i] < b.l[i];
    return false;
}
template <int F> static void put_fe_bytes(const Fe<F>& mont, uint8_t out[32]) {
    const Fe<F> c = fe_to_canonical<F>(mont);
    memcpy(out, c.l, 32);                                     // little-endian host
}
template <int F> static bool get_fe_bytes(const uint8_t in[32], Fe<F>& mont) {
    Fe<F> c;
    memcpy(c.l, in, 32);
    for (int i = 7; i >= 0; --i) {
        const u32 p = modulus_limb<F>(i);
        if (c.l[i] != p) { if (c.l[i] > p) return false; mont = fe_from_canonical<F>(c); return true; }
    }
    return false;                                             // equal to the modulus
}

template <int C> static void point_compress(const uint64_t xyz[12], uint8_t out[33]) {
    constexpr int B = CurveTraits<C>::BASE;
    Jac<C> j; memcpy(&j, xyz, 96);
    const Affine<C> a = jac_to_affine<C>(j);
    if (affine_is_identity<C>(a)) { memset(out, 0, 33); out[32] = 0x40; return; }
    put_fe_bytes<B>(a.x, out);
    const Fe<B> y = fe_to_canonical<B>(a.y), ny = fe_to_canonical<B>(fe_neg<B>(a.y));
    out[32] = canon_less<B>(ny, y) ? 0x80 : 0x00;             // y > -y: "negative"
}
template <int C> static bool point_decompress(const uint8_t in[33], uint64_t xyz[12]) {
    constexpr int B = CurveTraits<C>::BASE;
    const bool neg = (in[32] & 0x80) != 0, inf = (in[32] & 0x40) != 0;
    if (neg && inf) return false;
    Fe<B> x;
    if (!get_fe_bytes<B>(in, x)) return false;                // ark reads the field element before looking at the flags
    Jac<C> j;
    if (inf) { j.X = fe_one<B>(); j.Y = fe_one<B>(); j.Z = fe_zero<B>(); memcpy(xyz, &j, 96); return true; }
    const Fe<B> rhs = fe_add<B>(fe_mul<B>(fe_sqr<B>(x), x), fe_from_u64<B>(5));
    Fe<B> y;
    if (!fe_sqrt<B>(rhs, y)) return false;
    const Fe<B> ny = fe_neg<B>(y);
    const bool y_is_larger = canon_less<B>(fe_to_canonical<B>(ny), fe_to_canonical<B>(y));
    j.X = x; j.Y = (y_is_larger == neg) ? y : ny; j.Z = fe_one<B>();
    memcpy(xyz, &j, 96);
    return true;
}

static inline void put_u64(uint8_t* p, uint64_t v) { memcpy(p, &v, 8); }
static inline uint64_t get_u64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; }
static inline unsigned lg_of(uint64_t n) { unsigned l = 0; while (((uint64_t)1 << (l + 1)) <= n) ++l; return l; }

// flat instance (include/halo_b200.h): C[12] d[1] z[4] v[4] | Ls[12L] Rs[12L] U[12] c[4] hiding[1] C_bar[12] w_prime[4]
static size_t instance_bytes(unsigned L, bool hiding) { return 33 + 8 + 32 + 32 + 2 * (8 + 33 * (size_t)L) + 33 + 32 + (hiding ? 34 + 33 : 2); }

template <int C>
static int instance_serialize(const uint64_t* w, uint8_t* out, size_t cap, size_t* written) {
    constexpr int S = CurveTraits<C>::SCALAR;
    const uint64_t d = w[12];
    HALO_CHECK(d != ~(uint64_t)0 && ((d + 1) & d) == 0 && d + 1 <= ((uint64_t)1 << 40), HALO_EINVAL, "serialize: d + 1 is not a power of two");
    const unsigned L = lg_of(d + 1);
    const uint64_t* pr = w + 21;
    const bool hiding = pr[24 * L + 16] != 0;
    const size_t need = instance_bytes(L, hiding);
    if (written) *written = need;
    HALO_CHECK(out && cap >= need, HALO_EINVAL, "serialize: output buffer too small");
    uint8_t* p = out;
    auto scalar = [&](const uint64_t* s) { Fe<S> f; memcpy(&f, s, 32); put_fe_bytes<S>(f, p); p += 32; };
    point_compress<C>(w, p); p += 33;
    put_u64(p, d); p += 8;
    scalar(w + 13); scalar(w + 17);
    for (int half = 0; half < 2; ++half) {                    // Ls, then Rs
        put_u64(p, L); p += 8;
        for (unsigned i = 0; i < L; ++i) { point_compress<C>(pr + 12 * ((size_t)half * L + i), p); p += 33; }
    }
    point_compress<C>(pr + 24 * L, p); p += 33;
    scalar(pr + 24 * L + 12);
    *p++ = hiding ? 1 : 0;
    if (hiding) { point_compress<C>(pr + 24 * L + 17, p); p += 33; }
    *p++ = hiding ? 1 : 0;
    if (hiding) scalar(pr + 24 * L + 29);
    return HALO_OK;
}

template <int C>
static int instance_deserialize(const uint8_t* in, size_t len, uint64_t* w, size_t cap_words, size_t* consumed) {
    constexpr int S = CurveTraits<C>::SCALAR;
    size_t pos = 0;
    auto need = [&](size_t n) { return pos + n <= len; };
    HALO_CHECK(need(33 + 8), HALO_EINVAL, "deserialize: truncated");
    const uint64_t d = get_u64(in + 33);
    HALO_CHECK(d != ~(uint64_t)0 && ((d + 1) & d) == 0 && d + 1 <= ((uint64_t)1 << 40), HALO_EINVAL, "deserialize: d + 1 is not a power of two");
    const unsigned L = lg_of(d + 1);
    HALO_CHECK(cap_words >= 24 * (size_t)L + 54, HALO_EINVAL, "deserialize: output buffer too small");
    memset(w, 0, (24 * (size_t)L + 54) * 8);
    uint64_t* pr = w + 21;
    auto point = [&](uint64_t* dst) -> bool { if (!need(33) || !point_decompress<C>(in + pos, dst)) return false; pos += 33; return true; };
    auto scalar = [&](uint64_t* dst) -> bool { Fe<S> f; if (!need(32) || !get_fe_bytes<S>(in + pos, f)) return false; memcpy(dst, &f, 32); pos += 32; return true; };
    HALO_CHECK(point(w), HALO_EINVAL, "deserialize: C is not a valid compressed point");
    pos += 8; w[12] = d;
    HALO_CHECK(scalar(w + 13) && scalar(w + 17), HALO_EINVAL, "deserialize: z or v truncated or not below the modulus");
    for (int half = 0; half < 2; ++half) {
        HALO_CHECK(need(8) && get_u64(in + pos) == L, HALO_EINVAL, "deserialize: Ls / Rs must hold lg(d + 1) points");
        pos += 8;
        for (unsigned i = 0; i < L; ++i) HALO_CHECK(point(pr + 12 * ((size_t)half * L + i)), HALO_EINVAL, "deserialize: invalid point in Ls / Rs");
    }
    HALO_CHECK(point(pr + 24 * L), HALO_EINVAL, "deserialize: U is not a valid compressed point");
    HALO_CHECK(scalar(pr + 24 * L + 12), HALO_EINVAL, "deserialize: c truncated or not below the modulus");
    HALO_CHECK(need(1) && in[pos] <= 1, HALO_EINVAL, "deserialize: bad Option tag");
    const bool has_cbar = in[pos++] == 1;
    if (has_cbar) HALO_CHECK(point(pr + 24 * L + 17), HALO_EINVAL, "deserialize: C_bar is not a valid compressed point");
    HALO_CHECK(need(1) && in[pos] <= 1, HALO_EINVAL, "deserialize: bad Option tag");
    const bool has_w = in[pos++] == 1;
    if (has_w) HALO_CHECK(scalar(pr + 24 * L + 29), HALO_EINVAL, "deserialize: w' truncated or not below the modulus");
    HALO_CHECK(has_cbar == has_w, HALO_EINVAL, "deserialize: C_bar and w' must both be present or both absent (flat layout has one flag)");
    pr[24 * L + 16] = has_cbar ? 1 : 0;
    if (consumed) *consumed = pos;
    return HALO_OK;
}

}  // namespace halo

using namespace halo;
#define SER_TRY try {
#define SER_END } catch (...) { halo::last_err() = "serialize: unexpected exception"; return HALO_ECUDA; }

extern "C" {

int halo_point_compress(int curve, const uint64_t xyz[12], uint8_t out[33]) { SER_TRY
    HALO_CHECK((curve == 0 || curve == 1) && xyz && out, HALO_EINVAL, "point_compress: bad argument");
    if (curve == 0) point_compress<0>(xyz, out); else point_compress<1>(xyz, out);
    return HALO_OK;
SER_END }

int halo_point_decompress(int curve, const uint8_t in[33], uint64_t out_xyz[12]) { SER_TRY
    HALO_CHECK((curve == 0 || curve == 1) && in && out_xyz, HALO_EINVAL, "point_decompress: bad argument");
    const bool ok = curve == 0 ? point_decompress<0>(in, out_xyz) : point_decompress<1>(in, out_xyz);
    HALO_CHECK(ok, HALO_EINVAL, "point_decompress: not a valid compressed point (x >= p, x^3 + 5 not a square, or both flags set)");
    return HALO_OK;
SER_END }

int halo_scalar_serialize(int field, const uint64_t s[4], uint8_t out[32]) { SER_TRY
    HALO_CHECK((field == 0 || field == 1) && s && out, HALO_EINVAL, "scalar_serialize: bad argument");
    if (field == 0) { Fe<0> f; memcpy(&f, s, 32); put_fe_bytes<0>(f, out); } else { Fe<1> f; memcpy(&f, s, 32); put_fe_bytes<1>(f, out); }
    return HALO_OK;
SER_END }

int halo_scalar_deserialize(int field, const uint8_t in[32], uint64_t out[4]) { SER_TRY
    HALO_CHECK((field == 0 || field == 1) && in && out, HALO_EINVAL, "scalar_deserialize: bad argument");
    bool ok;
    if (field == 0) { Fe<0> f; ok = get_fe_bytes<0>(in, f); if (ok) memcpy(out, &f, 32); }
    else { Fe<1> f; ok = get_fe_bytes<1>(in, f); if (ok) memcpy(out, &f, 32); }
    HALO_CHECK(ok, HALO_EINVAL, "scalar_deserialize: not below the modulus");
    return HALO_OK;
SER_END }

int halo_instance_serialize(int curve, const uint64_t* instance, uint8_t* out, size_t cap, size_t* written) { SER_TRY
    HALO_CHECK((curve == 0 || curve == 1) && instance, HALO_EINVAL, "instance_serialize: bad argument");
    return curve == 0 ? instance_serialize<0>(instance, out, cap, written) : instance_serialize<1>(instance, out, cap, written);
SER_END }

int halo_instance_deserialize(int curve, const uint8_t* bytes, size_t len, uint64_t* out_instance, size_t cap_words, size_t* consumed) { SER_TRY
    HALO_CHECK((curve == 0 || curve == 1) && bytes && out_instance, HALO_EINVAL, "instance_deserialize: bad argument");
    return curve == 0 ? instance_deserialize<0>(bytes, len, out_instance, cap_words, consumed)
                      : instance_deserialize<1>(bytes, len, out_instance, cap_words, consumed);
SER_END }

}  // extern "C"
