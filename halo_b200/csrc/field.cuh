// 255-bit Montgomery arithmetic for the two Pasta fields on the sm_100a 32-bit integer pipe.
//
// Replaces, on the device, what the reference gets from ark-ff 0.5.0
// `Fp<MontBackend<_,4>,4>` (call sites: crates/group/src/group.rs:43-66,
// crates/accumulation/src/pcdl.rs:412-435 ...).  Same in-memory form as the reference
// (crates/group/src/wrappers.rs:603-616): Montgomery, R = 2^256, 4 x u64 LE limbs == 8 x u32 LE limbs.
//
// Both moduli have the shape  p = 1 + p1*2^32 + p2*2^64 + p3*2^96 + 2^254  with -p^{-1} = -1 (mod 2^32), so
//   * the Montgomery quotient digit is just  m = -t0  (no multiplication),
//   * m*p needs 3 wide multiplies (p1,p2,p3) instead of 8; the 2^254 term is a shift,
//   * p1+p2+p3 < 2^32, so a whole reduction column fits a 64-bit accumulator without carry flags.
// A modular multiplication is therefore 64 (product) + 24 (reduction) = 88 IMAD.WIDE on the FMA pipe
// (instead of 128 for a generic modulus) plus carry/shift work that rides the ALU pipe.
//
// The carry-chain primitives are individual PTX statements on the device.  With -DHALO_EMULATE_PTX the same
// algorithm compiles for the host with an emulated carry flag (used by tests/test_host_emulation to pin the
// device algorithm on a machine without a GPU).  Plain host builds (the host-side Fiat-Shamir transcript) use
// a 64-bit-limb implementation of the same functions.
#pragma once
#include <stdint.h>
#include <string.h>
#include <stdio.h>

#if defined(__CUDACC__)
#define HD __host__ __device__ __forceinline__
#define DEV __host__ __device__ __forceinline__
#else
#define HD inline
#define DEV inline
#endif
// Latency-bound kernels (a few warps walking a long dependent chain of point additions) want the multiplier as ONE
// shared subroutine: fully inlined, a point addition is ~70 KB of straight-line code and the instruction cache
// misses dominate.  Translation units built with -DHALO_NOINLINE_MUL get fe_mul / fe_sqr as real functions.
#if defined(__CUDACC__) && defined(HALO_NOINLINE_MUL)
#define MULDEV __host__ __device__ __noinline__
#else
#define MULDEV DEV
#endif

namespace halo {

typedef uint32_t u32;
typedef uint64_t u64;

// field id 0: halo Fp (ark_pallas::Fr), 1: halo Fq (ark_pallas::Fq)  -- crates/group/src/lib.rs:8-9
template <int FID> struct FieldConst;
template <> struct FieldConst<0> {
    static constexpr u32 P1 = 0x8c46eb21u, P2 = 0x0994a8ddu, P3 = 0x224698fcu;
};
template <> struct FieldConst<1> {
    static constexpr u32 P1 = 0x992d30edu, P2 = 0x094cf91bu, P3 = 0x224698fcu;
};
// limb k of the modulus
template <int FID> HD constexpr u32 modulus_limb(int k) {
    return k == 0 ? 1u : k == 1 ? FieldConst<FID>::P1 : k == 2 ? FieldConst<FID>::P2 : k == 3 ? FieldConst<FID>::P3
         : k == 7 ? 0x40000000u : 0u;
}

// ---------------------------------------------------------------------------------------------------------
// carry-chain primitives
// ---------------------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
#define HALO_PTX_PATH 1
DEV u32 add_cc(u32 a, u32 b) { u32 r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DEV u32 addc_cc(u32 a, u32 b) { u32 r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DEV u32 addc(u32 a, u32 b) { u32 r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DEV u32 sub_cc(u32 a, u32 b) { u32 r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DEV u32 subc_cc(u32 a, u32 b) { u32 r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DEV u32 subc(u32 a, u32 b) { u32 r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DEV u32 mad_lo_cc(u32 a, u32 b, u32 c) { u32 r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
DEV u32 madc_lo_cc(u32 a, u32 b, u32 c) { u32 r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
DEV u32 madc_hi_cc(u32 a, u32 b, u32 c) { u32 r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
DEV u32 madc_hi(u32 a, u32 b, u32 c) { u32 r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
#elif defined(HALO_EMULATE_PTX)
#define HALO_PTX_PATH 1
static thread_local u32 g_cc = 0;   // emulated carry / borrow flag
inline u32 add_cc(u32 a, u32 b) { u64 s = (u64)a + b; g_cc = (u32)(s >> 32); return (u32)s; }
inline u32 addc_cc(u32 a, u32 b) { u64 s = (u64)a + b + g_cc; g_cc = (u32)(s >> 32); return (u32)s; }
inline u32 addc(u32 a, u32 b) { u64 s = (u64)a + b + g_cc; return (u32)s; }
inline u32 sub_cc(u32 a, u32 b) { u64 s = (u64)a - b; g_cc = (u32)(s >> 63); return (u32)s; }
inline u32 subc_cc(u32 a, u32 b) { u64 s = (u64)a - b - g_cc; g_cc = (u32)(s >> 63); return (u32)s; }
inline u32 subc(u32 a, u32 b) { u64 s = (u64)a - b - g_cc; return (u32)s; }
inline u32 mad_lo_cc(u32 a, u32 b, u32 c) { u64 s = (u64)(u32)((u64)a * b) + c; g_cc = (u32)(s >> 32); return (u32)s; }
inline u32 madc_lo_cc(u32 a, u32 b, u32 c) { u64 s = (u64)(u32)((u64)a * b) + c + g_cc; g_cc = (u32)(s >> 32); return (u32)s; }
inline u32 madc_hi_cc(u32 a, u32 b, u32 c) { u64 s = (u64)(u32)(((u64)a * b) >> 32) + c + g_cc; g_cc = (u32)(s >> 32); return (u32)s; }
inline u32 madc_hi(u32 a, u32 b, u32 c) { u64 s = (u64)(u32)(((u64)a * b) >> 32) + c + g_cc; return (u32)s; }
#endif

// ---------------------------------------------------------------------------------------------------------
// element type
// ---------------------------------------------------------------------------------------------------------
template <int FID>
struct alignas(16) Fe {
    u32 l[8];
};

template <int FID> HD Fe<FID> fe_zero() { Fe<FID> r; for (int i = 0; i < 8; ++i) r.l[i] = 0; return r; }
template <int FID> HD bool fe_is_zero(const Fe<FID>& a) {
    u32 o = 0; for (int i = 0; i < 8; ++i) o |= a.l[i]; return o == 0;
}
template <int FID> HD bool fe_eq(const Fe<FID>& a, const Fe<FID>& b) {
    u32 o = 0; for (int i = 0; i < 8; ++i) o |= a.l[i] ^ b.l[i]; return o == 0;
}

#if defined(HALO_PTX_PATH)
// ---------------------------------------------------------------------------------------------------------
// device algorithm (also compiled for the host under HALO_EMULATE_PTX)
// ---------------------------------------------------------------------------------------------------------
template <int FID> DEV Fe<FID> fe_add(const Fe<FID>& a, const Fe<FID>& b) {
    u32 r[8], t[8];
    r[0] = add_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < 8; ++i) r[i] = addc_cc(a.l[i], b.l[i]);      // a+b < 2^256: no carry out
    t[0] = sub_cc(r[0], modulus_limb<FID>(0));
#pragma unroll
    for (int i = 1; i < 8; ++i) t[i] = subc_cc(r[i], modulus_limb<FID>(i));
    u32 borrow = subc(0, 0);                                          // 0xffffffff when r < p
    Fe<FID> o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.l[i] = borrow ? r[i] : t[i];
    return o;
}

template <int FID> DEV Fe<FID> fe_sub(const Fe<FID>& a, const Fe<FID>& b) {
    u32 r[8];
    r[0] = sub_cc(a.l[0], b.l[0]);
#pragma unroll
    for (int i = 1; i < 8; ++i) r[i] = subc_cc(a.l[i], b.l[i]);
    u32 mask = subc(0, 0);                                            // all ones when a < b
    Fe<FID> o;
    o.l[0] = add_cc(r[0], mask & modulus_limb<FID>(0));
#pragma unroll
    for (int i = 1; i < 8; ++i) o.l[i] = addc_cc(r[i], mask & modulus_limb<FID>(i));
    return o;
}

template <int FID> DEV Fe<FID> fe_neg(const Fe<FID>& a) { return fe_sub<FID>(fe_zero<FID>(), a); }
template <int FID> DEV Fe<FID> fe_dbl(const Fe<FID>& a) { return fe_add<FID>(a, a); }

// acc[off..off+8) += (x0,x1,x2,x3) * b with the four 64-bit products on consecutive limb pairs;
// the carry out is added into acc[off+8] when that limb exists.
template <int OFF, int N>
DEV void mad_row(u32 (&acc)[N], u32 x0, u32 x1, u32 x2, u32 x3, u32 b) {
    acc[OFF + 0] = mad_lo_cc(x0, b, acc[OFF + 0]);
    acc[OFF + 1] = madc_hi_cc(x0, b, acc[OFF + 1]);
    acc[OFF + 2] = madc_lo_cc(x1, b, acc[OFF + 2]);
    acc[OFF + 3] = madc_hi_cc(x1, b, acc[OFF + 3]);
    acc[OFF + 4] = madc_lo_cc(x2, b, acc[OFF + 4]);
    acc[OFF + 5] = madc_hi_cc(x2, b, acc[OFF + 5]);
    acc[OFF + 6] = madc_lo_cc(x3, b, acc[OFF + 6]);
    if (OFF + 8 < N) {
        acc[OFF + 7] = madc_hi_cc(x3, b, acc[OFF + 7]);
        acc[OFF + 8] = addc(acc[OFF + 8], 0);
    } else {
        acc[OFF + 7] = madc_hi(x3, b, acc[OFF + 7]);
    }
}

// Montgomery reduction of a 16-limb value t < p*2^256; returns t / 2^256 mod p, fully reduced.
// Column scan with a 64-bit accumulator: column j collects t[j] + m[j-1]*p1 + m[j-2]*p2 + m[j-3]*p3 (< 2^64
// because p1+p2+p3 < 2^32); m[j] = -column (mod 2^32).  The 2^254 term of p contributes M >> 2 to the upper
// half and (m[0] & 3) << 30 to column 7.
template <int FID> DEV Fe<FID> mont_reduce(const u32 (&t)[16]) {
    constexpr u32 P1 = FieldConst<FID>::P1, P2 = FieldConst<FID>::P2, P3 = FieldConst<FID>::P3;
    u32 m[8];
    u32 r[8];
    u64 acc = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        acc += t[j];
        if (j >= 1) acc += (u64)m[j - 1] * P1;
        if (j >= 2) acc += (u64)m[j - 2] * P2;
        if (j >= 3) acc += (u64)m[j - 3] * P3;
        if (j == 7) acc += (u64)((m[0] & 3u) << 30);
        m[j] = 0u - (u32)acc;
        acc += m[j];            // p0 = 1: cancels the low word
        acc >>= 32;
    }
#pragma unroll
    for (int j = 8; j < 16; ++j) {
        acc += t[j];
        if (j - 1 <= 7) acc += (u64)m[j - 1] * P1;
        if (j - 2 <= 7) acc += (u64)m[j - 2] * P2;
        if (j - 3 <= 7) acc += (u64)m[j - 3] * P3;
        r[j - 8] = (u32)acc;
        acc >>= 32;
    }
    // + (M >> 2)
    u32 s[8];
#pragma unroll
    for (int i = 0; i < 7; ++i) s[i] = (m[i] >> 2) | (m[i + 1] << 30);
    s[7] = m[7] >> 2;
    r[0] = add_cc(r[0], s[0]);
#pragma unroll
    for (int i = 1; i < 8; ++i) r[i] = addc_cc(r[i], s[i]);
    // result < 2p < 2^256: one conditional subtraction
    u32 d[8];
    d[0] = sub_cc(r[0], modulus_limb<FID>(0));
#pragma unroll
    for (int i = 1; i < 8; ++i) d[i] = subc_cc(r[i], modulus_limb<FID>(i));
    u32 borrow = subc(0, 0);
    Fe<FID> o;
#pragma unroll
    for (int i = 0; i < 8; ++i) o.l[i] = borrow ? r[i] : d[i];
    return o;
}

// 8x8 limb schoolbook product into even/odd-aligned accumulators so that every 64-bit partial product lands
// on an aligned register pair (one IMAD.WIDE.U32.X each).  ev holds columns 0..15, od holds columns 1..15.
template <int FID> MULDEV Fe<FID> fe_mul(const Fe<FID>& a, const Fe<FID>& b) {
    u32 ev[16], od[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { ev[i] = 0; od[i] = 0; }
    const u32 *x = a.l, *y = b.l;
    // row 0
    mad_row<0>(ev, x[0], x[2], x[4], x[6], y[0]);
    mad_row<0>(od, x[1], x[3], x[5], x[7], y[0]);
    // odd rows put a_even on od, a_odd on ev (shifted by 2); even rows the other way round
    mad_row<0>(od, x[0], x[2], x[4], x[6], y[1]);
    mad_row<2>(ev, x[1], x[3], x[5], x[7], y[1]);
    mad_row<2>(ev, x[0], x[2], x[4], x[6], y[2]);
    mad_row<2>(od, x[1], x[3], x[5], x[7], y[2]);
    mad_row<2>(od, x[0], x[2], x[4], x[6], y[3]);
    mad_row<4>(ev, x[1], x[3], x[5], x[7], y[3]);
    mad_row<4>(ev, x[0], x[2], x[4], x[6], y[4]);
    mad_row<4>(od, x[1], x[3], x[5], x[7], y[4]);
    mad_row<4>(od, x[0], x[2], x[4], x[6], y[5]);
    mad_row<6>(ev, x[1], x[3], x[5], x[7], y[5]);
    mad_row<6>(ev, x[0], x[2], x[4], x[6], y[6]);
    mad_row<6>(od, x[1], x[3], x[5], x[7], y[6]);
    mad_row<6>(od, x[0], x[2], x[4], x[6], y[7]);
    mad_row<8>(ev, x[1], x[3], x[5], x[7], y[7]);
    // t = ev + (od << 32)
    u32 t[16];
    t[0] = ev[0];
    t[1] = add_cc(ev[1], od[0]);
#pragma unroll
    for (int i = 2; i < 15; ++i) t[i] = addc_cc(ev[i], od[i - 1]);
    t[15] = addc(ev[15], od[14]);
    return mont_reduce<FID>(t);
}

// acc[OFF .. OFF+2L) += x * (a[J0], a[J0+2], ..., a[J0+2(L-1)]) on consecutive limb pairs; the carry out goes into
// acc[OFF+2L], which at that point holds at most a few earlier carries (see fe_sqr)
template <int OFF, int L, int J0, int N>
DEV void sqr_chain(u32 (&acc)[N], u32 x, const u32* a) {
    acc[OFF] = mad_lo_cc(x, a[J0], acc[OFF]);
    acc[OFF + 1] = madc_hi_cc(x, a[J0], acc[OFF + 1]);
#pragma unroll
    for (int t = 1; t < L; ++t) {
        acc[OFF + 2 * t] = madc_lo_cc(x, a[J0 + 2 * t], acc[OFF + 2 * t]);
        acc[OFF + 2 * t + 1] = madc_hi_cc(x, a[J0 + 2 * t], acc[OFF + 2 * t + 1]);
    }
    acc[OFF + 2 * L] = addc(acc[OFF + 2 * L], 0);
}

// Dedicated squaring: the 28 off-diagonal products a_i*a_j (i < j) once, doubled, plus the 8 diagonal squares:
// 36 wide multiplies instead of 64 (60 instead of 88 with the reduction).  Off-diagonal products with j - i odd
// land on odd columns (accumulator od, od[k] = column k+1), those with j - i even on even columns (ev).
template <int FID> MULDEV Fe<FID> fe_sqr(const Fe<FID>& a) {
    u32 ev[16], od[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { ev[i] = 0; od[i] = 0; }
    const u32* x = a.l;
    // j - i odd: row i starts at column 2i+1 = od[2i]
    sqr_chain<0, 4, 1>(od, x[0], x);     // a0 * (a1, a3, a5, a7)
    sqr_chain<2, 3, 2>(od, x[1], x);     // a1 * (a2, a4, a6)
    sqr_chain<4, 3, 3>(od, x[2], x);     // a2 * (a3, a5, a7)
    sqr_chain<6, 2, 4>(od, x[3], x);     // a3 * (a4, a6)
    sqr_chain<8, 2, 5>(od, x[4], x);     // a4 * (a5, a7)
    sqr_chain<10, 1, 6>(od, x[5], x);    // a5 * a6
    sqr_chain<12, 1, 7>(od, x[6], x);    // a6 * a7
    // j - i even: row i starts at column 2i+2 = ev[2i+2]
    sqr_chain<2, 3, 2>(ev, x[0], x);     // a0 * (a2, a4, a6)
    sqr_chain<4, 3, 3>(ev, x[1], x);     // a1 * (a3, a5, a7)
    sqr_chain<6, 2, 4>(ev, x[2], x);     // a2 * (a4, a6)
    sqr_chain<8, 2, 5>(ev, x[3], x);     // a3 * (a5, a7)
    sqr_chain<10, 1, 6>(ev, x[4], x);    // a4 * a6
    sqr_chain<12, 1, 7>(ev, x[5], x);    // a5 * a7
    // off = ev + (od << 32)
    u32 off[16];
    off[0] = ev[0];
    off[1] = add_cc(ev[1], od[0]);
#pragma unroll
    for (int i = 2; i < 15; ++i) off[i] = addc_cc(ev[i], od[i - 1]);
    off[15] = addc(ev[15], od[14]);
    // t = 2*off + sum_i a_i^2 * 2^(64 i): one carry chain over all 16 limbs
    u32 t[16];
#pragma unroll
    for (int i = 15; i >= 1; --i) t[i] = (off[i] << 1) | (off[i - 1] >> 31);
    t[0] = off[0] << 1;
    t[0] = mad_lo_cc(x[0], x[0], t[0]);
    t[1] = madc_hi_cc(x[0], x[0], t[1]);
#pragma unroll
    for (int i = 1; i < 7; ++i) {
        t[2 * i] = madc_lo_cc(x[i], x[i], t[2 * i]);
        t[2 * i + 1] = madc_hi_cc(x[i], x[i], t[2 * i + 1]);
    }
    t[14] = madc_lo_cc(x[7], x[7], t[14]);
    t[15] = madc_hi(x[7], x[7], t[15]);
    return mont_reduce<FID>(t);
}
#endif  // HALO_PTX_PATH

#if !defined(HALO_PTX_PATH)
// ---------------------------------------------------------------------------------------------------------
// plain host path (64-bit limbs): used by the host-side transcript / O(lg n) verifier work only
// ---------------------------------------------------------------------------------------------------------
namespace hostimpl {
typedef unsigned __int128 u128;
template <int FID> HD void modulus64(u64 p[4]) {
    p[0] = ((u64)FieldConst<FID>::P1 << 32) | 1u;
    p[1] = ((u64)FieldConst<FID>::P3 << 32) | FieldConst<FID>::P2;
    p[2] = 0; p[3] = 0x4000000000000000ULL;
}
// 64-bit accesses on both sides (little-endian host): a value stored as eight 32-bit words and reloaded as four 64-bit
// ones defeats store-to-load forwarding, which cost more than the multiplication itself in the serial sponge chain
HD void load(u64 w[4], const u32* l) { memcpy(w, l, 32); }
HD void store(u32* l, const u64 w[4]) { memcpy(l, w, 32); }
HD bool geq(const u64 a[4], const u64 b[4]) { for (int i = 3; i >= 0; --i) if (a[i] != b[i]) return a[i] > b[i]; return true; }
HD u64 add(u64 o[4], const u64 a[4], const u64 b[4]) { u128 c = 0; for (int i = 0; i < 4; ++i) { c += (u128)a[i] + b[i]; o[i] = (u64)c; c >>= 64; } return (u64)c; }
HD u64 sub(u64 o[4], const u64 a[4], const u64 b[4]) { u64 br = 0; for (int i = 0; i < 4; ++i) { u128 d = (u128)a[i] - b[i] - br; o[i] = (u64)d; br = (u64)(d >> 64) & 1; } return br; }
}  // namespace hostimpl

template <int FID> HD Fe<FID> fe_add(const Fe<FID>& a, const Fe<FID>& b) {
    using namespace hostimpl;
    u64 x[4], y[4], p[4]; load(x, a.l); load(y, b.l); modulus64<FID>(p);
    add(x, x, y); if (geq(x, p)) sub(x, x, p);
    Fe<FID> o; store(o.l, x); return o;
}
template <int FID> HD Fe<FID> fe_sub(const Fe<FID>& a, const Fe<FID>& b) {
    using namespace hostimpl;
    u64 x[4], y[4], p[4]; load(x, a.l); load(y, b.l); modulus64<FID>(p);
    if (sub(x, x, y)) add(x, x, p);
    Fe<FID> o; store(o.l, x); return o;
}
template <int FID> HD Fe<FID> fe_neg(const Fe<FID>& a) { return fe_sub<FID>(fe_zero<FID>(), a); }
template <int FID> HD Fe<FID> fe_dbl(const Fe<FID>& a) { return fe_add<FID>(a, a); }
namespace hostimpl {
// one Montgomery reduction step on 64-bit words: t = (t + m*p) >> 64 with m = -t[0]/p mod 2^64, for the moduli's
// shape p = [p0, p1, 0, 2^62]: two multiplications and a shift instead of four multiplications
template <int FID> HD void redc_step(u64* t /* 5 words + overflow in t[5] */) {
    const u64 p0 = ((u64)FieldConst<FID>::P1 << 32) | 1u, p1 = ((u64)FieldConst<FID>::P3 << 32) | FieldConst<FID>::P2;
    const u64 inv = (((u64)FieldConst<FID>::P1 - 1) << 32) | 0xffffffffULL;   // -p^{-1} mod 2^64
    const u64 m = t[0] * inv;
    u128 c = ((u128)m * p0 + t[0]) >> 64;
    c += (u128)m * p1 + t[1]; t[0] = (u64)c; c >>= 64;
    c += t[2]; t[1] = (u64)c; c >>= 64;
    c += ((u128)m << 62) + t[3]; t[2] = (u64)c; c >>= 64;
    c += t[4]; t[3] = (u64)c; t[4] = t[5] + (u64)(c >> 64);
}
// w[0..8) += x * y (schoolbook); returns the carry out of word 7
HD u64 mac_wide(u64 w[8], const u64 x[4], const u64 y[4]) {
    u64 carry_out = 0;
    for (int i = 0; i < 4; ++i) {
        u128 c = 0;
        for (int j = 0; j < 4; ++j) { c += (u128)x[j] * y[i] + w[i + j]; w[i + j] = (u64)c; c >>= 64; }
        for (int k = i + 4; k < 8 && c; ++k) { c += w[k]; w[k] = (u64)c; c >>= 64; }
        carry_out += (u64)c;
    }
    return carry_out;
}
// Montgomery reduction of a 512-bit value T < 3 p^2 (so that the result is below 2p before the final subtraction)
template <int FID> HD void redc_wide(u64 out[4], const u64 w[8]) {
    u64 p[4]; modulus64<FID>(p);
    u64 t[6] = {w[0], w[1], w[2], w[3], 0, 0};
    for (int i = 0; i < 4; ++i) {
        // bring in the next high word, then divide by 2^64
        u128 c = (u128)t[4] + w[4 + i];
        t[4] = (u64)c; t[5] = (u64)(c >> 64);
        redc_step<FID>(t);
    }
    if (t[4] || geq(t, p)) sub(t, t, p);
    for (int i = 0; i < 4; ++i) out[i] = t[i];
}
}  // namespace hostimpl

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__CUDA_ARCH__) && !defined(HALO_NO_HOST_ADX)
#define HALO_HOST_ADX 1
namespace hostimpl {
// BMI2 + ADX path of the host multiplication (the serial Fiat-Shamir sponge is ~5 k dependent multiplications per
// challenge and bounds AS_DL over many instances): the interleaved (CIOS) Montgomery product, one row per asm block.
// A row adds x * y_i to the 5-word accumulator and divides by 2^64 after adding m p (m = -t0 / p mod 2^64); mulx leaves
// the flags alone, adcx / adox run the low-word and high-word carry chains side by side.  Every output is early-clobber:
// the accumulator words are written before the last x_j is read, and an input with the same VALUE (x_j = 0 and a zero
// accumulator word, x_0 = y_i in a squaring) would otherwise be given the same register.  K = {p0, p1, p3, -p^-1 mod
// 2^64} (p2 = 0).  Bounds: t < 2p before a row, t + x y_i + m p < 2^64 * 2p < 2^320, so five words never overflow.
inline bool host_has_adx() {
    static const bool ok = __builtin_cpu_supports("bmi2") && __builtin_cpu_supports("adx");
    return ok;
}
#define HALO_MONT_ROW(T0, T1, T2, T3, T4, YI)                                                                         \
    {                                                                                                                 \
        unsigned long long d_ = (YI), lo_, hi_, z_;                                                                   \
        asm("xorl %k[z], %k[z]\n\t"                                                                                   \
            "mulx %[x0], %[lo], %[hi]\n\t adox %[lo], %[t0]\n\t adcx %[hi], %[t1]\n\t"                                \
            "mulx %[x1], %[lo], %[hi]\n\t adox %[lo], %[t1]\n\t adcx %[hi], %[t2]\n\t"                                \
            "mulx %[x2], %[lo], %[hi]\n\t adox %[lo], %[t2]\n\t adcx %[hi], %[t3]\n\t"                                \
            "mulx %[x3], %[lo], %[hi]\n\t adox %[lo], %[t3]\n\t adcx %[hi], %[t4]\n\t"                                \
            "adox %[z], %[t4]\n\t"                                                                                    \
            "mov %[t0], %%rdx\n\t imulq 24(%[k]), %%rdx\n\t"                                                          \
            "xorl %k[z], %k[z]\n\t"                                                                                   \
            "mulx 0(%[k]), %[lo], %[hi]\n\t adox %[lo], %[t0]\n\t adcx %[hi], %[t1]\n\t"                              \
            "mulx 8(%[k]), %[lo], %[hi]\n\t adox %[lo], %[t1]\n\t adcx %[hi], %[t2]\n\t"                              \
            "adox %[z], %[t2]\n\t"                                                                                    \
            "mulx 16(%[k]), %[lo], %[hi]\n\t adox %[lo], %[t3]\n\t adcx %[z], %[t3]\n\t"                              \
            "adox %[hi], %[t4]\n\t adcx %[z], %[t4]\n\t"                                                              \
            : [t0] "+&r"(T0), [t1] "+&r"(T1), [t2] "+&r"(T2), [t3] "+&r"(T3), [t4] "+&r"(T4), "+&d"(d_), [lo] "=&r"(lo_), \
              [hi] "=&r"(hi_), [z] "=&r"(z_)                                                                          \
            : [x0] "rm"(x0), [x1] "rm"(x1), [x2] "rm"(x2), [x3] "rm"(x3), [k] "r"(K)                                   \
            : "cc");                                                                                                  \
    }
template <int FID> inline void mont_mul_adx(u64 r[4], const u64 x[4], const u64 y[4]) {
    typedef unsigned long long ull;
    static const ull K[4] = {((ull)FieldConst<FID>::P1 << 32) | 1u, ((ull)FieldConst<FID>::P3 << 32) | FieldConst<FID>::P2,
                             0x4000000000000000ULL, (((ull)FieldConst<FID>::P1 - 1) << 32) | 0xffffffffULL};
    const ull x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
    ull t0 = 0, t1 = 0, t2 = 0, t3 = 0, t4 = 0;
    HALO_MONT_ROW(t0, t1, t2, t3, t4, y[0]); t0 = 0;      // the low word is zero after a row: it becomes the next top word
    HALO_MONT_ROW(t1, t2, t3, t4, t0, y[1]); t1 = 0;
    HALO_MONT_ROW(t2, t3, t4, t0, t1, y[2]); t2 = 0;
    HALO_MONT_ROW(t3, t4, t0, t1, t2, y[3]);
    // (t4, t0, t1, t2) < 2p: one conditional subtraction
    ull s0, s1, s2, s3;
    unsigned char bw;
    bw = __builtin_ia32_sbb_u64(0, t4, K[0], &s0);
    bw = __builtin_ia32_sbb_u64(bw, t0, K[1], &s1);
    bw = __builtin_ia32_sbb_u64(bw, t1, 0, &s2);
    bw = __builtin_ia32_sbb_u64(bw, t2, K[2], &s3);
    r[0] = bw ? t4 : s0; r[1] = bw ? t0 : s1; r[2] = bw ? t1 : s2; r[3] = bw ? t2 : s3;
}
#undef HALO_MONT_ROW
// The three-term dot product of the sponge's MDS rows with ONE reduction: per word i of the second operands three
// multiply-accumulate rows into a six-word accumulator, then one reduction row.  The result (T + M p) / 2^256 with
// T < 3 p^2 is below 2p, as in redc_wide.
#define HALO_MAC_ROW(T0, T1, T2, T3, T4, T5, X, YI)                                                                   \
    {                                                                                                                 \
        unsigned long long d_ = (YI), lo_, hi_, z_;                                                                   \
        asm("xorl %k[z], %k[z]\n\t"                                                                                   \
            "mulx %[x0], %[lo], %[hi]\n\t adox %[lo], %[t0]\n\t adcx %[hi], %[t1]\n\t"                                \
            "mulx %[x1], %[lo], %[hi]\n\t adox %[lo], %[t1]\n\t adcx %[hi], %[t2]\n\t"                                \
            "mulx %[x2], %[lo], %[hi]\n\t adox %[lo], %[t2]\n\t adcx %[hi], %[t3]\n\t"                                \
            "mulx %[x3], %[lo], %[hi]\n\t adox %[lo], %[t3]\n\t adcx %[hi], %[t4]\n\t"                                \
            "adox %[z], %[t4]\n\t adcx %[z], %[t5]\n\t adox %[z], %[t5]\n\t"                                          \
            : [t0] "+&r"(T0), [t1] "+&r"(T1), [t2] "+&r"(T2), [t3] "+&r"(T3), [t4] "+&r"(T4), [t5] "+&r"(T5),         \
              [lo] "=&r"(lo_), [hi] "=&r"(hi_), [z] "=&r"(z_)                                                         \
            : "d"(d_), [x0] "rm"(X[0]), [x1] "rm"(X[1]), [x2] "rm"(X[2]), [x3] "rm"(X[3])                             \
            : "cc");                                                                                                  \
    }
#define HALO_REDC_ROW(T0, T1, T2, T3, T4, T5)                                                                         \
    {                                                                                                                 \
        unsigned long long d_, lo_, hi_, z_;                                                                          \
        asm("mov %[t0], %%rdx\n\t imulq 24(%[k]), %%rdx\n\t"                                                          \
            "xorl %k[z], %k[z]\n\t"                                                                                   \
            "mulx 0(%[k]), %[lo], %[hi]\n\t adox %[lo], %[t0]\n\t adcx %[hi], %[t1]\n\t"                              \
            "mulx 8(%[k]), %[lo], %[hi]\n\t adox %[lo], %[t1]\n\t adcx %[hi], %[t2]\n\t"                              \
            "adox %[z], %[t2]\n\t"                                                                                    \
            "mulx 16(%[k]), %[lo], %[hi]\n\t adox %[lo], %[t3]\n\t adcx %[z], %[t3]\n\t"                              \
            "adox %[hi], %[t4]\n\t adcx %[z], %[t4]\n\t"                                                              \
            "adox %[z], %[t5]\n\t adcx %[z], %[t5]\n\t"                                                               \
            : [t0] "+&r"(T0), [t1] "+&r"(T1), [t2] "+&r"(T2), [t3] "+&r"(T3), [t4] "+&r"(T4), [t5] "+&r"(T5),         \
              "=&d"(d_), [lo] "=&r"(lo_), [hi] "=&r"(hi_), [z] "=&r"(z_)                                              \
            : [k] "r"(K)                                                                                              \
            : "cc");                                                                                                  \
    }
template <int FID> inline void mont_dot3_adx(u64 r[4], const u64 a0[4], const u64 b0[4], const u64 a1[4], const u64 b1[4],
                                             const u64 a2[4], const u64 b2[4]) {
    typedef unsigned long long ull;
    static const ull K[4] = {((ull)FieldConst<FID>::P1 << 32) | 1u, ((ull)FieldConst<FID>::P3 << 32) | FieldConst<FID>::P2,
                             0x4000000000000000ULL, (((ull)FieldConst<FID>::P1 - 1) << 32) | 0xffffffffULL};
    ull t0 = 0, t1 = 0, t2 = 0, t3 = 0, t4 = 0, t5 = 0;
#define HALO_DOT_STEP(T0, T1, T2, T3, T4, T5, I)                                                                      \
    HALO_MAC_ROW(T0, T1, T2, T3, T4, T5, a0, b0[I]) HALO_MAC_ROW(T0, T1, T2, T3, T4, T5, a1, b1[I])                   \
    HALO_MAC_ROW(T0, T1, T2, T3, T4, T5, a2, b2[I]) HALO_REDC_ROW(T0, T1, T2, T3, T4, T5)
    HALO_DOT_STEP(t0, t1, t2, t3, t4, t5, 0) t0 = 0;      // the low word is zero after the reduction row: the next top word
    HALO_DOT_STEP(t1, t2, t3, t4, t5, t0, 1) t1 = 0;
    HALO_DOT_STEP(t2, t3, t4, t5, t0, t1, 2) t2 = 0;
    HALO_DOT_STEP(t3, t4, t5, t0, t1, t2, 3)
#undef HALO_DOT_STEP
    // (t4, t5, t0, t1) with top words t2 = 0: below 2p, one conditional subtraction
    ull s0, s1, s2, s3;
    unsigned char bw;
    bw = __builtin_ia32_sbb_u64(0, t4, K[0], &s0);
    bw = __builtin_ia32_sbb_u64(bw, t5, K[1], &s1);
    bw = __builtin_ia32_sbb_u64(bw, t0, 0, &s2);
    bw = __builtin_ia32_sbb_u64(bw, t1, K[2], &s3);
    r[0] = bw ? t4 : s0; r[1] = bw ? t5 : s1; r[2] = bw ? t0 : s2; r[3] = bw ? t1 : s3;
}
#undef HALO_MAC_ROW
#undef HALO_REDC_ROW
}  // namespace hostimpl
#endif

template <int FID> HD Fe<FID> fe_mul(const Fe<FID>& a, const Fe<FID>& b) {
    using namespace hostimpl;
    u64 x[4], y[4], p[4]; load(x, a.l); load(y, b.l); modulus64<FID>(p);
#if defined(HALO_HOST_ADX)
    // the five-word bound of the ADX rows needs both operands below 2^254 (< p); callers that reduce a raw 256-bit integer
    // by multiplying it with R^2 (fe_from_canonical on a value of the other field) take the generic path with its sixth word
    if (((x[3] | y[3]) >> 62) == 0 && host_has_adx()) {
        u64 r[4];
        mont_mul_adx<FID>(r, x, y);
#ifdef HALO_ADX_SELFCHECK
        {
            u64 t[6] = {0, 0, 0, 0, 0, 0};
            for (int i = 0; i < 4; ++i) {
                u128 c = 0;
                for (int j = 0; j < 4; ++j) { c += (u128)x[j] * y[i] + t[j]; t[j] = (u64)c; c >>= 64; }
                c += t[4]; t[4] = (u64)c; t[5] = (u64)(c >> 64);
                redc_step<FID>(t);
            }
            if (t[4] || geq(t, p)) sub(t, t, p);
            if (t[0] != r[0] || t[1] != r[1] || t[2] != r[2] || t[3] != r[3])
                fprintf(stderr, "ADX MISMATCH F%d x=%016llx %016llx %016llx %016llx y=%016llx %016llx %016llx %016llx adx=%016llx.. gen=%016llx..\n", FID,
                        (unsigned long long)x[3], (unsigned long long)x[2], (unsigned long long)x[1], (unsigned long long)x[0],
                        (unsigned long long)y[3], (unsigned long long)y[2], (unsigned long long)y[1], (unsigned long long)y[0],
                        (unsigned long long)r[3], (unsigned long long)t[3]);
        }
#endif
        Fe<FID> o; store(o.l, r); return o;
    }
#endif
    u64 t[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < 4; ++i) {
        u128 c = 0;
        for (int j = 0; j < 4; ++j) { c += (u128)x[j] * y[i] + t[j]; t[j] = (u64)c; c >>= 64; }
        c += t[4]; t[4] = (u64)c; t[5] = (u64)(c >> 64);
        redc_step<FID>(t);
    }
    if (t[4] || geq(t, p)) sub(t, t, p);
    Fe<FID> o; store(o.l, t); return o;
}
// a0*b0 + a1*b1 + a2*b2 with one reduction (the MDS rows of the host sponge)
template <int FID> HD Fe<FID> fe_dot3(const Fe<FID>& a0, const Fe<FID>& b0, const Fe<FID>& a1, const Fe<FID>& b1,
                                      const Fe<FID>& a2, const Fe<FID>& b2) {
    using namespace hostimpl;
#if defined(HALO_HOST_ADX)
    {
        u64 xa[4], ya[4], xb[4], yb[4], xc[4], yc[4];
        load(xa, a0.l); load(ya, b0.l); load(xb, a1.l); load(yb, b1.l); load(xc, a2.l); load(yc, b2.l);
        if (((xa[3] | ya[3] | xb[3] | yb[3] | xc[3] | yc[3]) >> 62) == 0 && host_has_adx()) {      // all below 2^254 < p
            u64 r[4];
            mont_dot3_adx<FID>(r, xa, ya, xb, yb, xc, yc);
            Fe<FID> o; store(o.l, r); return o;
        }
    }
#endif
    u64 x[4], y[4], w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    load(x, a0.l); load(y, b0.l); mac_wide(w, x, y);
    load(x, a1.l); load(y, b1.l); mac_wide(w, x, y);
    load(x, a2.l); load(y, b2.l); mac_wide(w, x, y);          // 3 p^2 < 2^512: no carry out
    u64 r[4]; redc_wide<FID>(r, w);
    Fe<FID> o; store(o.l, r); return o;
}
template <int FID> HD Fe<FID> fe_sqr(const Fe<FID>& a) { return fe_mul<FID>(a, a); }
#endif  // !HALO_PTX_PATH

#if defined(HALO_PTX_PATH)
template <int FID> DEV Fe<FID> fe_dot3(const Fe<FID>& a0, const Fe<FID>& b0, const Fe<FID>& a1, const Fe<FID>& b1,
                                       const Fe<FID>& a2, const Fe<FID>& b2) {
    return fe_add<FID>(fe_add<FID>(fe_mul<FID>(a0, b0), fe_mul<FID>(a1, b1)), fe_mul<FID>(a2, b2));
}
#endif

// ---------------------------------------------------------------------------------------------------------
// constants and derived operations (all paths)
// ---------------------------------------------------------------------------------------------------------
template <int FID> HD Fe<FID> fe_one() {       // R mod p
    Fe<FID> r;
    r.l[0] = 0xfffffffdu;
    r.l[1] = FID == 0 ? 0x5b2b3e9cu : 0x34786d38u;
    r.l[2] = FID == 0 ? 0xe3420567u : 0xe41914adu;
    r.l[3] = 0x992c350bu;
    r.l[4] = 0xffffffffu; r.l[5] = 0xffffffffu; r.l[6] = 0xffffffffu; r.l[7] = 0x3fffffffu;
    return r;
}
template <int FID> HD Fe<FID> fe_r2() {        // R^2 mod p
    Fe<FID> r;
    if (FID == 0) {
        r.l[0] = 0x0000000fu; r.l[1] = 0xfc9678ffu; r.l[2] = 0x891a16e3u; r.l[3] = 0x67bb433du;
        r.l[4] = 0x04ccf590u; r.l[5] = 0x7fae2310u; r.l[6] = 0x7ccfdaa9u; r.l[7] = 0x096d41afu;
    } else {
        r.l[0] = 0x0000000fu; r.l[1] = 0x8c78ecb3u; r.l[2] = 0x8b0de0e7u; r.l[3] = 0xd7d30dbdu;
        r.l[4] = 0xc3c95d18u; r.l[5] = 0x7797a99bu; r.l[6] = 0x7b9cb714u; r.l[7] = 0x096d41afu;
    }
    return r;
}
// canonical integer (8 x u32, < p) -> Montgomery
template <int FID> HD Fe<FID> fe_from_canonical(const Fe<FID>& c) { return fe_mul<FID>(c, fe_r2<FID>()); }
// Montgomery -> canonical integer
template <int FID> HD Fe<FID> fe_to_canonical(const Fe<FID>& a) {
    Fe<FID> one = fe_zero<FID>(); one.l[0] = 1;
    return fe_mul<FID>(a, one);
}
template <int FID> HD Fe<FID> fe_from_u64(u64 x) {
    Fe<FID> c = fe_zero<FID>(); c.l[0] = (u32)x; c.l[1] = (u32)(x >> 32);
    return fe_from_canonical<FID>(c);
}
// a^e, e given as 8 x u32 little-endian
template <int FID> HD Fe<FID> fe_pow(const Fe<FID>& a, const u32* e, int nbits = 256) {
    Fe<FID> acc = fe_one<FID>();
    for (int i = nbits - 1; i >= 0; --i) {
        acc = fe_sqr<FID>(acc);
        if ((e[i >> 5] >> (i & 31)) & 1) acc = fe_mul<FID>(acc, a);
    }
    return acc;
}
template <int FID> HD Fe<FID> fe_inv_safegcd(const Fe<FID>& a);
// Inverse; 0 -> 0.  Device: Fermat ladder.  Plain host path (transcript, verifier arithmetic: an inversion per challenge
// and per point brought to affine form): the divstep inversion further down, a third of the ladder's time on a host core.
template <int FID> HD Fe<FID> fe_inv(const Fe<FID>& a) {
#if !defined(HALO_PTX_PATH)
    return fe_inv_safegcd<FID>(a);
#endif
    u32 e[8];
    for (int i = 0; i < 8; ++i) e[i] = modulus_limb<FID>(i);
    e[0] = 0xffffffffu; e[1] -= 1;            // p - 2  (p0 = 1 -> borrow from limb 1)
    return fe_pow<FID>(a, e, 255);
}

// ---------------------------------------------------------------------------------------------------------
// Inversion by Bernstein-Yang divsteps ("safegcd") on signed 30-bit limbs, constant time: 20 x 30 divsteps, each
// group of 30 steps computed on the low words only and applied to the full-size numbers as a 2x2 matrix (the
// layout of libsecp256k1's modinv32; its proof of the 590-step bound for 256-bit inputs covers these 255-bit moduli).
// ~1800 wide multiplies and ~13 k ALU operations: about a fifth of the Fermat ladder (254 squarings + ~60
// multiplications), and the ALU work rides the pipe the multiplier leaves idle.  Branch-free, so the lanes of a warp
// stay converged.  Both moduli are 1 mod 2^32, so modulus^-1 mod 2^30 = 1.
// Input and output are Montgomery representatives: for X = a R the routine returns R^2 / X = a^-1 R; 0 -> 0.
// ---------------------------------------------------------------------------------------------------------
template <int FID> HD constexpr u32 modulus_limb30(int i) {
    const int bit = 30 * i, w = bit >> 5, sh = bit & 31;
    const u64 lo = w < 8 ? modulus_limb<FID>(w) : 0u, hi = w + 1 < 8 ? modulus_limb<FID>(w + 1) : 0u;
    return (u32)((((hi << 32) | lo) >> sh) & 0x3fffffffu);
}
struct Signed30 { int32_t v[9]; };

template <int FID> HD Signed30 fe_to_signed30(const Fe<FID>& a) {
    Signed30 r;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 9; ++i) {
        const int bit = 30 * i, w = bit >> 5, sh = bit & 31;
        const u64 lo = w < 8 ? a.l[w] : 0u, hi = w + 1 < 8 ? a.l[w + 1] : 0u;
        r.v[i] = (int32_t)((((hi << 32) | lo) >> sh) & 0x3fffffffu);
    }
    return r;
}
// limbs in [0, 2^30), value < 2^256
template <int FID> HD Fe<FID> fe_from_signed30(const Signed30& s) {
    Fe<FID> r;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int w = 0; w < 8; ++w) {
        const int bit = 32 * w, i = bit / 30, sh = bit % 30;
        u64 acc = (u64)(u32)s.v[i] >> sh;
        if (i + 1 < 9) acc |= (u64)(u32)s.v[i + 1] << (30 - sh);
        if (i + 2 < 9) acc |= (u64)(u32)s.v[i + 2] << (60 - sh);
        r.l[w] = (u32)acc;
    }
    return r;
}

struct Trans2x2 { int32_t u, v, q, r; };

// 30 divsteps on the low 30+ bits of f (odd) and g; the matrix t satisfies t * [f, g] = 2^30 * [f', g']
HD int32_t divsteps_30(int32_t zeta, u32 f0, u32 g0, Trans2x2& t) {
    u32 u = 1, v = 0, q = 0, r = 1, f = f0, g = g0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 30; ++i) {
        u32 mask1 = (u32)(zeta >> 31);                 // zeta < 0
        const u32 mask2 = 0u - (g & 1u);               // g odd
        const u32 x = (f ^ mask1) - mask1, y = (u ^ mask1) - mask1, z = (v ^ mask1) - mask1;   // +-(f, u, v)
        g += x & mask2; q += y & mask2; r += z & mask2;
        mask1 &= mask2;
        zeta = (int32_t)((u32)zeta ^ mask1) - 1;       // -zeta - 2  or  zeta - 1
        f += g & mask1; u += q & mask1; v += r & mask1;
        g >>= 1; u <<= 1; v <<= 1;
    }
    t.u = (int32_t)u; t.v = (int32_t)v; t.q = (int32_t)q; t.r = (int32_t)r;
    return zeta;
}
// (f, g) <- t * (f, g) / 2^30  (exact)
HD void update_fg_30(Signed30& f, Signed30& g, const Trans2x2& t) {
    const int32_t M30 = 0x3fffffff;
    int64_t cf = (int64_t)t.u * f.v[0] + (int64_t)t.v * g.v[0];
    int64_t cg = (int64_t)t.q * f.v[0] + (int64_t)t.r * g.v[0];
    cf >>= 30; cg >>= 30;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 1; i < 9; ++i) {
        const int32_t fi = f.v[i], gi = g.v[i];
        cf += (int64_t)t.u * fi + (int64_t)t.v * gi;
        cg += (int64_t)t.q * fi + (int64_t)t.r * gi;
        f.v[i - 1] = (int32_t)cf & M30; cf >>= 30;
        g.v[i - 1] = (int32_t)cg & M30; cg >>= 30;
    }
    f.v[8] = (int32_t)cf; g.v[8] = (int32_t)cg;
}
// (d, e) <- t * (d, e) / 2^30 mod p, kept in (-2p, p)
template <int FID> HD void update_de_30(Signed30& d, Signed30& e, const Trans2x2& t) {
    const int32_t M30 = 0x3fffffff;
    const int32_t sd = d.v[8] >> 31, se = e.v[8] >> 31;
    int32_t md = (t.u & sd) + (t.v & se), me = (t.q & sd) + (t.r & se);
    int64_t cd = (int64_t)t.u * d.v[0] + (int64_t)t.v * e.v[0];
    int64_t ce = (int64_t)t.q * d.v[0] + (int64_t)t.r * e.v[0];
    md -= (int32_t)(((u32)cd + (u32)md) & (u32)M30);   // modulus^-1 mod 2^30 = 1
    me -= (int32_t)(((u32)ce + (u32)me) & (u32)M30);
    cd += (int64_t)(int32_t)modulus_limb30<FID>(0) * md;
    ce += (int64_t)(int32_t)modulus_limb30<FID>(0) * me;
    cd >>= 30; ce >>= 30;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 1; i < 9; ++i) {
        const int32_t di = d.v[i], ei = e.v[i];
        cd += (int64_t)t.u * di + (int64_t)t.v * ei;
        ce += (int64_t)t.q * di + (int64_t)t.r * ei;
        if (modulus_limb30<FID>(i) != 0u) {            // limbs 5..7 of both moduli are zero
            cd += (int64_t)(int32_t)modulus_limb30<FID>(i) * md;
            ce += (int64_t)(int32_t)modulus_limb30<FID>(i) * me;
        }
        d.v[i - 1] = (int32_t)cd & M30; cd >>= 30;
        e.v[i - 1] = (int32_t)ce & M30; ce >>= 30;
    }
    d.v[8] = (int32_t)cd; e.v[8] = (int32_t)ce;
}
// r in (-2p, p) -> [0, p), negated first when sign < 0
template <int FID> HD void normalize_30(Signed30& r, int32_t sign) {
    const int32_t M30 = 0x3fffffff;
    int32_t cond_add = r.v[8] >> 31;
    const int32_t cond_neg = sign >> 31;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 9; ++i) {
        r.v[i] += (int32_t)modulus_limb30<FID>(i) & cond_add;
        r.v[i] = (r.v[i] ^ cond_neg) - cond_neg;
    }
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 8; ++i) { r.v[i + 1] += r.v[i] >> 30; r.v[i] &= M30; }
    cond_add = r.v[8] >> 31;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 9; ++i) r.v[i] += (int32_t)modulus_limb30<FID>(i) & cond_add;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < 8; ++i) { r.v[i + 1] += r.v[i] >> 30; r.v[i] &= M30; }
}

template <int FID> HD Fe<FID> fe_inv_safegcd(const Fe<FID>& a) {
    Signed30 d, e = fe_to_signed30<FID>(fe_r2<FID>()), f, g = fe_to_signed30<FID>(a);
    for (int i = 0; i < 9; ++i) { d.v[i] = 0; f.v[i] = (int32_t)modulus_limb30<FID>(i); }
    int32_t zeta = -1;
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (int it = 0; it < 20; ++it) {
        Trans2x2 t;
        zeta = divsteps_30(zeta, (u32)f.v[0] | ((u32)f.v[1] << 30), (u32)g.v[0] | ((u32)g.v[1] << 30), t);
        update_de_30<FID>(d, e, t);
        update_fg_30(f, g, t);
    }
    normalize_30<FID>(d, f.v[8]);                      // f = +-1 (or +-p when a = 0, where d = 0 anyway)
    return fe_from_signed30<FID>(d);
}

}  // namespace halo
