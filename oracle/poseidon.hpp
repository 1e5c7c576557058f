// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Restatement of the reference's Poseidon sponge and Fiat-Shamir wrapper:
//   crates/poseidon/src/inner_sponge.rs:3-115  (x^7 s-box, 55 full rounds,
//       width 3 = rate 2 + capacity 1, round = sbox -> MDS -> + round constants)
//   crates/poseidon/src/outer_sponge.rs:16-99  (protocol label, point / scalar
//       absorption across the two fields of the cycle, challenge squeeze)
// Constants: crates/group/src/poseidon_consts.rs:17-609 (data, extracted by
// tools/extract_reference_data.py).  Pinned by the Kimchi KATs (G4).
#pragma once
#include "curve.hpp"

namespace orc {

#include "poseidon_consts.inc"

template <int FID>
struct PoseidonSponge {
    typedef Fe<FID> F;
    F st[3];
    bool squeezing;   // false: Absorbed(cnt), true: Squeezed(cnt)
    int cnt;

    PoseidonSponge() { reset(); }
    void reset() { st[0] = st[1] = st[2] = F::zero(); squeezing = false; cnt = 0; }

    static F sbox(F x) {
        F sq = x.sqr();
        x = x * sq;
        sq = sq.sqr();
        return x * sq;
    }
    void permute() {
        for (int r = 0; r < 55; ++r) {
            F s[3];
            for (int i = 0; i < 3; ++i) s[i] = sbox(st[i]);
            for (int i = 0; i < 3; ++i) {
                F acc = F::zero();
                for (int j = 0; j < 3; ++j)
                    acc = acc + s[j] * F::from_raw((const u64*)POSEIDON_MDS[FID][i * 3 + j]);
                st[i] = acc + F::from_raw((const u64*)POSEIDON_RC[FID][r * 3 + i]);
            }
        }
    }
    void absorb(const F& x) {
        if (!squeezing && cnt < 2) { st[cnt] = st[cnt] + x; cnt++; }
        else if (!squeezing) { permute(); cnt = 1; st[0] = st[0] + x; }
        else { squeezing = false; cnt = 1; st[0] = st[0] + x; }
    }
    F squeeze() {
        if (squeezing && cnt < 2) { return st[cnt++]; }
        permute(); squeezing = true; cnt = 1;
        return st[0];
    }
};

enum Protocol { PROTO_PCDL = 0, PROTO_ASDL = 1, PROTO_PLONK = 2, PROTO_SIGNATURE = 3 };

// Sponge<P>: state lives in P's base field, challenges land in P's scalar field.
template <int CID>
struct Transcript {
    enum { BASE = CurveTraits<CID>::BASE, SCALAR = CurveTraits<CID>::SCALAR };
    typedef Fe<BASE> BF;
    typedef Fe<SCALAR> SF;
    PoseidonSponge<BASE> sponge;

    // modulus(halo Fp, id 0) > modulus(halo Fq, id 1)
    static bool scalar_lt_base() { return !geq4(field_params(SCALAR).p, field_params(BASE).p); }

    explicit Transcript(Protocol label) { sponge.absorb(BF::from_u64((u64)label)); }

    void absorb_affine(const Affine<CID>& g) {
        if (g.inf) { sponge.absorb(BF::zero()); sponge.absorb(BF::zero()); }
        else { sponge.absorb(g.x); sponge.absorb(g.y); }
    }
    void absorb_g(const Jac<CID>& g) { absorb_affine(g.to_affine()); }
    void absorb_fq(const BF& x) { sponge.absorb(x); }
    void absorb_fr(const SF& x) {
        u64 w[4]; x.to_canonical(w);
        if (scalar_lt_base()) {
            sponge.absorb(BF::from_canonical(w));
        } else {
            u64 low = w[0] & 1;
            u64 h[4];
            for (int i = 0; i < 4; ++i) h[i] = (w[i] >> 1) | (i < 3 ? (w[i + 1] << 63) : 0);
            sponge.absorb(BF::from_canonical(h));
            sponge.absorb(low ? BF::one() : BF::zero());
        }
    }
    SF challenge() {
        u64 w[4]; sponge.squeeze().to_canonical(w);
        if (scalar_lt_base()) {
            u64 h[4];
            for (int i = 0; i < 4; ++i) h[i] = (w[i] >> 1) | (i < 3 ? (w[i + 1] << 63) : 0);
            return SF::from_canonical(h);
        }
        return SF::from_canonical(w);
    }
};

}  // namespace orc
