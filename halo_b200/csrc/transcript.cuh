// Host-side Fiat-Shamir transcript of the product: the Poseidon sponge and its cross-field wrapper.
//
// Stands in for crates/poseidon/src/inner_sponge.rs:3-115 (x^7 s-box, 55 full rounds of sbox -> MDS -> +rc, width
// 3 = rate 2 + capacity 1) and crates/poseidon/src/outer_sponge.rs:16-99 (protocol label; points absorbed as
// affine (x, y) with the identity as (0, 0); scalars and challenges carried between the two fields of the cycle).
// It stays on the host: the IPA needs one challenge per round from two freshly produced points, a serial chain
// of ~4 permutations (~5 k field multiplications) that a GPU thread would run slower than the host core.
// Constants are data extracted from crates/group/src/poseidon_consts.rs (poseidon_consts.inc).
#pragma once
#include "curve.cuh"
#include <cstring>

namespace halo {

namespace poseidon_data {
#include "poseidon_consts.inc"
}

enum TranscriptLabel { LABEL_PCDL = 0, LABEL_ASDL = 1, LABEL_PLONK = 2, LABEL_SIGNATURE = 3 };

template <int F> inline Fe<F> fe_from_words(const unsigned long long* w) { Fe<F> r; memcpy(&r, w, 32); return r; }

// shift a 256-bit little-endian integer (8 x u32) right by one bit
inline void shr1(u32* v) {
    for (int i = 0; i < 8; ++i) v[i] = (v[i] >> 1) | (i < 7 ? (v[i + 1] << 31) : 0u);
}

// Sponge over field F.  `filled` counts rate cells written since the last permutation while absorbing, or cells
// already handed out while squeezing.
template <int F>
class PoseidonHost {
public:
    PoseidonHost() { clear(); }
    void clear() { s_[0] = s_[1] = s_[2] = fe_zero<F>(); absorbing_ = true; filled_ = 0; }

    void absorb(const Fe<F>& x) {
        if (!absorbing_) { absorbing_ = true; filled_ = 0; }          // Squeezed(_) -> Absorbed(1) without permuting
        else if (filled_ == 2) { permute(); filled_ = 0; }
        s_[filled_] = fe_add<F>(s_[filled_], x);
        ++filled_;
    }
    Fe<F> squeeze() {
        if (absorbing_ || filled_ == 2) { permute(); absorbing_ = false; filled_ = 0; }
        return s_[filled_++];
    }

private:
    static Fe<F> pow7(const Fe<F>& x) {
        Fe<F> x2 = fe_sqr<F>(x), x3 = fe_mul<F>(x, x2), x4 = fe_sqr<F>(x2);
        return fe_mul<F>(x3, x4);
    }
    void permute() {
#if defined(HALO_HOST_ADX)
        if (hostimpl::host_has_adx()) { permute_adx(); return; }
#endif
        for (int r = 0; r < 55; ++r) {
            Fe<F> t[3] = {pow7(s_[0]), pow7(s_[1]), pow7(s_[2])};
            for (int i = 0; i < 3; ++i) {
                const unsigned long long (*M)[4] = poseidon_data::POSEIDON_MDS[F] + 3 * i;
                s_[i] = fe_add<F>(fe_from_words<F>(poseidon_data::POSEIDON_RC[F][3 * r + i]),
                                  fe_dot3<F>(t[0], fe_from_words<F>(M[0]), t[1], fe_from_words<F>(M[1]), t[2], fe_from_words<F>(M[2])));
            }
        }
    }
#if defined(HALO_HOST_ADX)
    // The same 55 rounds on 64-bit limbs throughout (BMI2 + ADX rows of field.cuh): the state, the s-box intermediates and
    // the MDS rows never pass through the 32-bit-limb element type.  Every value stays below p (products and sums are
    // fully reduced), which is what the five-word rows need.
    void permute_adx() {
        using namespace hostimpl;
        u64 st[3][4], t[3][4], p[4];
        modulus64<F>(p);
        for (int i = 0; i < 3; ++i) load(st[i], s_[i].l);
        for (int r = 0; r < 55; ++r) {
            for (int i = 0; i < 3; ++i) {
                u64 x2[4], x3[4], x4[4];
                mont_mul_adx<F>(x2, st[i], st[i]);
                mont_mul_adx<F>(x4, x2, x2);
                mont_mul_adx<F>(x3, x2, st[i]);
                mont_mul_adx<F>(t[i], x4, x3);
            }
            for (int i = 0; i < 3; ++i) {
                const unsigned long long (*M)[4] = poseidon_data::POSEIDON_MDS[F] + 3 * i;
                u64 d[4];
                mont_dot3_adx<F>(d, t[0], (const u64*)M[0], t[1], (const u64*)M[1], t[2], (const u64*)M[2]);
                add(d, d, (const u64*)poseidon_data::POSEIDON_RC[F][3 * r + i]);      // both below p < 2^255: no carry out
                if (geq(d, p)) sub(d, d, p);
                for (int k = 0; k < 4; ++k) st[i][k] = d[k];
            }
        }
        for (int i = 0; i < 3; ++i) store(s_[i].l, st[i]);
    }
#endif
    Fe<F> s_[3];
    bool absorbing_;
    int filled_;
};

// Sponge<P>: the state lives in curve C's base field, challenges land in its scalar field.
template <int C>
class Transcript {
    static constexpr int B = CurveTraits<C>::BASE, S = CurveTraits<C>::SCALAR;
    // halo Fp (id 0) has the larger modulus: Pallas scalars exceed the Pallas base field, Vesta's do not
    static constexpr bool SCALAR_FITS = (S == 1);

public:
    explicit Transcript(TranscriptLabel label) { sponge_.absorb(fe_from_u64<B>((u64)label)); }

    void absorb_affine(const Affine<C>& g) { sponge_.absorb(g.x); sponge_.absorb(g.y); }   // identity is (0,0)
    void absorb_point(const XYZZ<C>& g) { absorb_affine(xyzz_to_affine<C>(g)); }
    void absorb_base(const Fe<B>& x) { sponge_.absorb(x); }
    void absorb_scalar(const Fe<S>& x) {
        Fe<S> c = fe_to_canonical<S>(x);
        Fe<B> v;
        if (SCALAR_FITS) {
            memcpy(v.l, c.l, 32);
            sponge_.absorb(fe_from_canonical<B>(v));
        } else {                                   // high bits, then the low bit (outer_sponge.rs:73-83)
            u32 low = c.l[0] & 1u;
            shr1(c.l);
            memcpy(v.l, c.l, 32);
            sponge_.absorb(fe_from_canonical<B>(v));
            sponge_.absorb(low ? fe_one<B>() : fe_zero<B>());
        }
    }
    Fe<S> challenge() {
        Fe<B> c = fe_to_canonical<B>(sponge_.squeeze());
        if (SCALAR_FITS) shr1(c.l);                // base value may exceed the scalar modulus: drop the low bit
        Fe<S> v;
        memcpy(v.l, c.l, 32);
        return fe_from_canonical<S>(v);
    }

private:
    PoseidonHost<B> sponge_;
};

}  // namespace halo
