// Experimental 255-bit Montgomery multipliers, compared against field.cuh's fe_mul by tools/mulbench/mulbench.cu.
#pragma once
#include "field.cuh"
namespace halo {
#if defined(HALO_PTX_PATH)

struct ColAcc { u32 lo, hi, ov; };
DEV void col_mac0(ColAcc& p, u32 a, u32 b) {      // first product of a column pair: p.hi == 0, p.lo small => no overflow
    p.lo = mad_lo_cc(a, b, p.lo);
    p.hi = madc_hi(a, b, p.hi);
}
DEV void col_mac(ColAcc& p, u32 a, u32 b) {
    p.lo = mad_lo_cc(a, b, p.lo);
    p.hi = madc_hi_cc(a, b, p.hi);
    p.ov = addc(p.ov, 0);
}

// product scanning: 64-bit column-pair accumulators on two alignments, overflow counted per pair and fed
// into the next pair of the same alignment; no carry chain lives longer than one product.
template <int FID> DEV void mul_wide_comba(const u32* x, const u32* y, u32 (&t)[16]) {
    u32 elo[8], ehi[8], olo[8], ohi[8];
    u32 carry_e = 0, carry_o = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        ColAcc p; p.lo = carry_e; p.hi = 0; p.ov = 0;
        bool first = true;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int j = 2 * k - i;
            if (j < 0 || j > 7) continue;
            if (first) { col_mac0(p, x[i], y[j]); first = false; } else col_mac(p, x[i], y[j]);
        }
        elo[k] = p.lo; ehi[k] = p.hi; carry_e = p.ov;
    }
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        ColAcc p; p.lo = carry_o; p.hi = 0; p.ov = 0;
        bool first = true;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int j = 2 * k + 1 - i;
            if (j < 0 || j > 7) continue;
            if (first) { col_mac0(p, x[i], y[j]); first = false; } else col_mac(p, x[i], y[j]);
        }
        olo[k] = p.lo; ohi[k] = p.hi; carry_o = p.ov;
    }
    olo[7] = carry_o; ohi[7] = 0;
    t[0] = elo[0];
    t[1] = add_cc(ehi[0], olo[0]);
#pragma unroll
    for (int i = 1; i < 8; ++i) {
        t[2 * i] = addc_cc(elo[i], ohi[i - 1]);
        if (i < 7) t[2 * i + 1] = addc_cc(ehi[i], olo[i]);
        else t[2 * i + 1] = addc(ehi[i], olo[i]);
    }
}

template <int FID> DEV Fe<FID> fe_mul_comba(const Fe<FID>& a, const Fe<FID>& b) {
    u32 t[16];
    mul_wide_comba<FID>(a.l, b.l, t);
    return mont_reduce<FID>(t);
}


// Variant 2: the merge of the two alignments and the Montgomery column scan in one pass, written with 64-bit
// integers so that ptxas can use three-input adds with two carry predicates.
template <int FID> DEV Fe<FID> fe_mul_comba2(const Fe<FID>& a, const Fe<FID>& b) {
    constexpr u32 P1 = FieldConst<FID>::P1, P2 = FieldConst<FID>::P2, P3 = FieldConst<FID>::P3;
    const u32 *x = a.l, *y = b.l;
    u32 ew[16], ow[16];     // ew[j]: word of the even-aligned pairs in column j; ow[j]: word of the odd-aligned pairs
    u32 carry_e = 0, carry_o = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        ColAcc p; p.lo = carry_e; p.hi = 0; p.ov = 0;
        bool first = true;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int j = 2 * k - i;
            if (j < 0 || j > 7) continue;
            if (first) { col_mac0(p, x[i], y[j]); first = false; } else col_mac(p, x[i], y[j]);
        }
        ew[2 * k] = p.lo; ew[2 * k + 1] = p.hi; carry_e = p.ov;
    }
    ow[0] = 0;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        ColAcc p; p.lo = carry_o; p.hi = 0; p.ov = 0;
        bool first = true;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int j = 2 * k + 1 - i;
            if (j < 0 || j > 7) continue;
            if (first) { col_mac0(p, x[i], y[j]); first = false; } else col_mac(p, x[i], y[j]);
        }
        ow[2 * k + 1] = p.lo; ow[2 * k + 2] = p.hi; carry_o = p.ov;
    }
    ow[15] = carry_o;
    u32 m[8], r[8];
    u64 acc = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        acc += (u64)ew[j] + ow[j];
        if (j >= 1) acc += (u64)m[j - 1] * P1;
        if (j >= 2) acc += (u64)m[j - 2] * P2;
        if (j >= 3) acc += (u64)m[j - 3] * P3;
        if (j == 7) acc += (u64)((m[0] & 3u) << 30);
        u32 lo = (u32)acc;
        m[j] = 0u - lo;
        acc = (acc >> 32) + (lo != 0 ? 1u : 0u);
    }
#pragma unroll
    for (int j = 8; j < 16; ++j) {
        u32 s = j < 15 ? ((m[j - 8] >> 2) | (m[j - 7] << 30)) : (m[7] >> 2);
        acc += (u64)ew[j] + ow[j];
        acc += s;
        if (j - 1 <= 7) acc += (u64)m[j - 1] * P1;
        if (j - 2 <= 7) acc += (u64)m[j - 2] * P2;
        if (j - 3 <= 7) acc += (u64)m[j - 3] * P3;
        r[j - 8] = (u32)acc;
        acc >>= 32;
    }
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
#endif  // HALO_PTX_PATH
}  // namespace halo
