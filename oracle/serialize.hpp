// ORACLE — test infrastructure.  CPU restatement of the ark-serialize 0.5.0 compressed encodings the reference derives
// for its PC_DL / AS_DL objects:
//   crates/accumulation/src/pcdl.rs:26-34    #[derive(CanonicalSerialize, CanonicalDeserialize)] Instance
//   crates/accumulation/src/pcdl.rs:159-168  ... EvalProof { Ls, Rs, U, c, C_bar: Option, w_prime: Option }
//   crates/accumulation/src/acc.rs:22-26     ... Accumulator { q }
//   crates/accumulation/src/main.rs:59-91    qs.bin = Vec<(usize, Instance, Accumulator)>::serialize_compressed
// The encoders live in un-vendored crates (ark-serialize / ark-ff / ark-ec 0.5.0); their published behaviour:
//   ark-ff  Fp::serialize_with_flags     canonical integer, little-endian, ceil((255 + flag bits) / 8) bytes, flags OR-ed
//                                        into the top bits of the last byte; deserialisation fails for values >= p
//   ark-ec  SWFlags                      bit 7 = YIsNegative (y > -y), bit 6 = PointAtInfinity, both = invalid
//   ark-ec  Affine::serialize compressed x with flags (33 bytes here); Projective serialises its affine form
//   ark-ff  SqrtPrecomputation::TonelliShanks (two-adicity 32, non-residue 5)
//   ark-serialize  usize -> u64 LE; Vec<T> -> u64 LE length + items; Option<T> -> bool byte + item; bool -> 0 / 1
// PINNED by tests/test_serialize.py against the bytes of the reference's own crates/accumulation/.precompute/qs.bin
// (written by an older struct layout whose point / scalar / container encodings are the same).
#pragma once
#include "pcdl.hpp"
#include <cstdint>
#include <vector>

namespace orc {
namespace ser {

template <int FID>
bool sqrt_tonelli_shanks(const Fe<FID>& elem, Fe<FID>* out) {
    typedef Fe<FID> F;
    if (elem.is_zero()) { *out = F::zero(); return true; }
    F z = root_of_unity<FID>(32);                             // QUADRATIC_NONRESIDUE_TO_TRACE = 5^t
    u64 e[4], p1[4]; u64 one[4] = {1, 0, 0, 0};
    sub4(p1, F::P().p, one);                                  // TRACE_MINUS_ONE_DIV_TWO = ((p - 1) >> 32 - 1) / 2 = (p - 1) >> 33
    for (int i = 0; i < 4; ++i) {
        const unsigned __int128 pair = ((unsigned __int128)(i + 1 < 4 ? p1[i + 1] : 0) << 64) | p1[i];
        e[i] = (u64)(pair >> 33);
    }
    F w = elem.pow(e);
    F x = w * elem, b = x * w;
    unsigned v = 32;
    while (!(b == F::one())) {
        unsigned k = 0;
        F b2k = b;
        while (!(b2k == F::one())) { b2k = b2k.sqr(); ++k; if (k == 32) return false; }
        if (k >= v) return false;
        const unsigned j = v - k;
        w = z;
        for (unsigned i = 1; i < j; ++i) w = w.sqr();
        z = w.sqr(); b = b * z; x = x * w; v = k;
    }
    if (!(x.sqr() == elem)) return false;
    *out = x;
    return true;
}

static inline bool less4(const u64 a[4], const u64 b[4]) {
    for (int i = 3; i >= 0; --i) if (a[i] != b[i]) return a[i] < b[i];
    return false;
}

template <int FID> void write_fe(const Fe<FID>& f, std::vector<uint8_t>& out) {
    u64 w[4]; f.to_canonical(w);
    for (int i = 0; i < 4; ++i) for (int b = 0; b < 8; ++b) out.push_back((uint8_t)(w[i] >> (8 * b)));
}
template <int FID> bool read_fe(const uint8_t* in, Fe<FID>* f) {
    u64 w[4] = {0, 0, 0, 0};
    for (int i = 0; i < 4; ++i) for (int b = 0; b < 8; ++b) w[i] |= (u64)in[8 * i + b] << (8 * b);
    if (!less4(w, Fe<FID>::P().p)) return false;
    *f = Fe<FID>::from_canonical(w);
    return true;
}
static inline void write_u64(u64 v, std::vector<uint8_t>& out) { for (int b = 0; b < 8; ++b) out.push_back((uint8_t)(v >> (8 * b))); }
static inline u64 read_u64(const uint8_t* in) { u64 v = 0; for (int b = 0; b < 8; ++b) v |= (u64)in[b] << (8 * b); return v; }

template <int CID> void write_point(const Jac<CID>& p, std::vector<uint8_t>& out) {
    typedef Fe<CurveTraits<CID>::BASE> F;
    const Affine<CID> a = p.to_affine();
    if (a.inf) { for (int i = 0; i < 32; ++i) out.push_back(0); out.push_back(0x40); return; }
    write_fe(a.x, out);
    u64 y[4], ny[4];
    a.y.to_canonical(y); a.y.neg().to_canonical(ny);
    out.push_back(less4(ny, y) ? 0x80 : 0x00);                 // SWFlags::from_y_coordinate: y <= -y is "positive"
    (void)sizeof(F);
}
template <int CID> bool read_point(const uint8_t* in, Jac<CID>* p) {
    typedef Fe<CurveTraits<CID>::BASE> F;
    const bool neg = (in[32] >> 7) & 1, inf = (in[32] >> 6) & 1;
    if (neg && inf) return false;
    F x;
    if (!read_fe(in, &x)) return false;
    if (inf) { *p = Jac<CID>::identity(); return true; }
    F y;
    if (!sqrt_tonelli_shanks<CurveTraits<CID>::BASE>(x.sqr() * x + F::from_u64(5), &y)) return false;
    u64 yc[4], nyc[4];
    y.to_canonical(yc); y.neg().to_canonical(nyc);
    const F smaller = less4(nyc, yc) ? y.neg() : y, larger = less4(nyc, yc) ? y : y.neg();
    Affine<CID> a; a.x = x; a.y = neg ? larger : smaller; a.inf = false;
    *p = Jac<CID>::from_affine(a);
    return true;
}

template <int CID> void write_instance(const Instance<CID>& q, std::vector<uint8_t>& out) {
    write_point(q.C, out);
    write_u64((u64)q.d, out);
    write_fe(q.z, out); write_fe(q.v, out);
    write_u64(q.pi.Ls.size(), out); for (const auto& L : q.pi.Ls) write_point(L, out);
    write_u64(q.pi.Rs.size(), out); for (const auto& R : q.pi.Rs) write_point(R, out);
    write_point(q.pi.U, out);
    write_fe(q.pi.c, out);
    out.push_back(q.pi.hiding ? 1 : 0); if (q.pi.hiding) write_point(q.pi.C_bar, out);
    out.push_back(q.pi.hiding ? 1 : 0); if (q.pi.hiding) write_fe(q.pi.w_prime, out);
}
// returns bytes consumed, or 0 on malformed input
template <int CID> size_t read_instance(const uint8_t* in, size_t len, Instance<CID>* q) {
    size_t pos = 0;
    auto need = [&](size_t n) { return pos + n <= len; };
    if (!need(33) || !read_point(in + pos, &q->C)) return 0;
    pos += 33;
    if (!need(8)) return 0;
    q->d = (size_t)read_u64(in + pos);
    pos += 8;
    if (!need(64) || !read_fe(in + pos, &q->z) || !read_fe(in + pos + 32, &q->v)) return 0;
    pos += 64;
    for (int half = 0; half < 2; ++half) {
        if (!need(8)) return 0;
        const u64 cnt = read_u64(in + pos); pos += 8;
        if (cnt > 64) return 0;
        auto& vec = half ? q->pi.Rs : q->pi.Ls;
        vec.resize(cnt);
        for (u64 i = 0; i < cnt; ++i) { if (!need(33) || !read_point(in + pos, &vec[i])) return 0;
    pos += 33; }
    }
    if (!need(33) || !read_point(in + pos, &q->pi.U)) return 0;
    pos += 33;
    if (!need(32) || !read_fe(in + pos, &q->pi.c)) return 0;
    pos += 32;
    if (!need(1) || in[pos] > 1) return 0;
    const bool has_c = in[pos++] == 1;
    q->pi.C_bar = Jac<CID>::identity();
    if (has_c) { if (!need(33) || !read_point(in + pos, &q->pi.C_bar)) return 0;
    pos += 33; }
    if (!need(1) || in[pos] > 1) return 0;
    const bool has_w = in[pos++] == 1;
    q->pi.w_prime = Fe<CurveTraits<CID>::SCALAR>::zero();
    if (has_w) { if (!need(32) || !read_fe(in + pos, &q->pi.w_prime)) return 0;
    pos += 32; }
    if (has_c != has_w) return 0;                              // the flat FFI layout carries one flag for both Options
    q->pi.hiding = has_c;
    return pos;
}

}  // namespace ser
}  // namespace orc
