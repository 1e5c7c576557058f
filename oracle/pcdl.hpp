// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Restatement of the reference's Pedersen commitment, PC_DL and AS_DL:
//   crates/accumulation/src/pedersen.rs:7-27      commit
//   crates/accumulation/src/pcdl.rs:191-233       HPoly::{get_poly,eval}
//   crates/accumulation/src/pcdl.rs:275-287       commit
//   crates/accumulation/src/pcdl.rs:326-473       open_without_eval / open
//   crates/accumulation/src/pcdl.rs:483-554       succinct_check
//   crates/accumulation/src/pcdl.rs:563-583       check
//   crates/accumulation/src/acc.rs:71-110,128-217 common_subroutine / prover / verifier / decider
//   crates/group/src/group.rs:43-66               scalar_dot / point_dot / construct_powers
//   crates/group/src/pp.rs:10-94                  PublicParams
// Pinned by the genesis accumulators at crates/plonk/src/frontend/ivc/mod.rs:196-292.
#pragma once
#include "msm.hpp"
#include "ntt.hpp"
#include "poseidon.hpp"
#include <stdexcept>
#include <string>

namespace orc {

template <int CID>
struct PublicParams {
    std::vector<Affine<CID>> Gs;
    Jac<CID> S, H;
    bool loaded = false;
    size_t D() const { return Gs.size() - 1; }
};
template <int CID> PublicParams<CID>& pp() { static PublicParams<CID> p; return p; }

extern unsigned g_threads;

template <class Fn>
static void par_for_idx(size_t n, const Fn& fn) {
    par_for(n, n >= 64 ? g_threads : 1, [&](size_t lo, size_t hi) { for (size_t j = lo; j < hi; ++j) fn(j); });
}

template <int CID>
struct EvalProof {
    typedef Fe<CurveTraits<CID>::SCALAR> SF;
    std::vector<Jac<CID>> Ls, Rs;
    Jac<CID> U;
    SF c;
    bool hiding = false;
    Jac<CID> C_bar;
    SF w_prime;
};

template <int CID>
struct Instance {
    typedef Fe<CurveTraits<CID>::SCALAR> SF;
    Jac<CID> C;
    size_t d;
    SF z, v;
    EvalProof<CID> pi;
};

// flat FFI layout shared with the product's C ABI (include/halo_b200.h):
// C[12] d[1] z[4] v[4] Ls[12L] Rs[12L] U[12] c[4] hiding[1] C_bar[12] w_prime[4]
static inline size_t instance_words(unsigned lg_n) { return 54 + 24 * (size_t)lg_n; }

template <int CID>
void instance_to_flat(const Instance<CID>& q, u64* w) {
    unsigned L = (unsigned)q.pi.Ls.size();
    q.C.to_raw(w); w += 12;
    *w++ = q.d;
    q.z.to_raw(w); w += 4;
    q.v.to_raw(w); w += 4;
    for (unsigned i = 0; i < L; ++i) { q.pi.Ls[i].to_raw(w); w += 12; }
    for (unsigned i = 0; i < L; ++i) { q.pi.Rs[i].to_raw(w); w += 12; }
    q.pi.U.to_raw(w); w += 12;
    q.pi.c.to_raw(w); w += 4;
    *w++ = q.pi.hiding ? 1 : 0;
    if (q.pi.hiding) { q.pi.C_bar.to_raw(w); w += 12; q.pi.w_prime.to_raw(w); w += 4; }
    else { memset(w, 0, 16 * 8); }
}

template <int CID>
Instance<CID> instance_from_flat(const u64* w) {
    typedef Fe<CurveTraits<CID>::SCALAR> SF;
    Instance<CID> q;
    q.C = Jac<CID>::from_raw(w); w += 12;
    q.d = *w++;
    unsigned L = 0; { size_t n = q.d + 1; while (((size_t)1 << L) < n) ++L; }
    q.z = SF::from_raw(w); w += 4;
    q.v = SF::from_raw(w); w += 4;
    q.pi.Ls.resize(L); q.pi.Rs.resize(L);
    for (unsigned i = 0; i < L; ++i) { q.pi.Ls[i] = Jac<CID>::from_raw(w); w += 12; }
    for (unsigned i = 0; i < L; ++i) { q.pi.Rs[i] = Jac<CID>::from_raw(w); w += 12; }
    q.pi.U = Jac<CID>::from_raw(w); w += 12;
    q.pi.c = SF::from_raw(w); w += 4;
    q.pi.hiding = *w++ != 0;
    q.pi.C_bar = Jac<CID>::from_raw(w); w += 12;
    q.pi.w_prime = SF::from_raw(w);
    return q;
}

// ---- group.rs helpers ----
template <class SF>
SF scalar_dot(const SF* a, const SF* b, size_t n) {
    SF acc = SF::zero();
    for (size_t i = 0; i < n; ++i) acc = acc + a[i] * b[i];
    return acc;
}
template <class SF>
std::vector<SF> construct_powers(const SF& z, size_t n) {
    std::vector<SF> zs(n);
    SF cur = SF::one();
    for (size_t i = 0; i < n; ++i) { zs[i] = cur; cur = cur * z; }
    return zs;
}
template <int CID>
Jac<CID> point_dot_affine(const typename Jac<CID>::S* xs, size_t nx, const Affine<CID>* gs, size_t ng) {
    size_t k = nx < ng ? nx : ng;
    return msm<CID>(gs, (const u64*)xs, k, g_threads);
}

// ---- pedersen.rs:7-27 ----
template <int CID>
Jac<CID> pedersen_commit(const typename Jac<CID>::S* w, const Affine<CID>* gs, size_t ng,
                         const typename Jac<CID>::S* ms, size_t nm) {
    if (ng < nm) throw std::runtime_error("ms must be larger than Gs");
    Jac<CID> acc = msm<CID>(gs, (const u64*)ms, nm, g_threads);
    if (w) return pp<CID>().S.mul(*w).add(acc);
    return acc;
}

// trailing zero coefficients are stripped, as DensePolynomial does
template <class SF>
size_t poly_len(const SF* coeffs, size_t n) {
    while (n > 0 && coeffs[n - 1].is_zero()) --n;
    return n;
}

// ---- pcdl.rs:275-287 ----
template <int CID>
Jac<CID> pcdl_commit(const typename Jac<CID>::S* coeffs, size_t len, size_t d, const typename Jac<CID>::S* w) {
    auto& P = pp<CID>();
    size_t n = d + 1;
    len = poly_len(coeffs, len);
    if (n & (n - 1)) throw std::runtime_error("n is not a power of two");
    if (len > n) throw std::runtime_error("p_deg <= d");
    if (d > P.D()) throw std::runtime_error("d <= D");
    return pedersen_commit<CID>(w, P.Gs.data(), n, coeffs, len);
}

// ---- pcdl.rs:191-233 ----
template <int CID>
struct HPoly {
    typedef Fe<CurveTraits<CID>::SCALAR> SF;
    std::vector<SF> xis;
    std::vector<SF> get_poly() const {
        size_t lg_n = xis.size() - 1;
        std::vector<SF> h(1, SF::one());
        h.reserve((size_t)1 << lg_n);
        for (size_t i = 0; i < lg_n; ++i) {
            size_t cur = h.size();
            const SF& xi = xis[lg_n - i];
            for (size_t j = 0; j < cur; ++j) h.push_back(h[j] * xi);
        }
        return h;
    }
    SF eval(const SF& z) const {
        size_t lg_n = xis.size() - 1;
        SF one = SF::one();
        SF v = one + xis[lg_n] * z;
        SF zi = z;
        for (size_t i = 1; i < lg_n; ++i) { zi = zi.sqr(); v = v * (one + xis[lg_n - i] * zi); }
        return v;
    }
};

template <class SF>
SF poly_eval(const SF* coeffs, size_t len, const SF& z) {
    SF acc = SF::zero();
    for (size_t i = len; i-- > 0;) acc = acc * z + coeffs[i];
    return acc;
}

// ---- pcdl.rs:326-453.  Hiding randomness (q of degree d-1, w_bar) is supplied by the caller. ----
template <int CID>
EvalProof<CID> pcdl_open_without_eval(const typename Jac<CID>::S* coeffs, size_t len, const Jac<CID>& C, size_t d,
                                     const typename Jac<CID>::S& z, const typename Jac<CID>::S& v,
                                     const typename Jac<CID>::S* w,
                                     const typename Jac<CID>::S* hide_q, const typename Jac<CID>::S* hide_wbar) {
    typedef typename Jac<CID>::S SF;
    auto& P = pp<CID>();
    Transcript<CID> tr(PROTO_PCDL);
    size_t n = d + 1;
    if (n <= 1 || (n & (n - 1))) throw std::runtime_error("n is not a power of two");
    len = poly_len(coeffs, len);
    if (len > n) throw std::runtime_error("p.degree() <= d");
    if (d > P.D()) throw std::runtime_error("d <= pp.D");
    unsigned lg_n = 0; while (((size_t)1 << lg_n) < n) ++lg_n;

    EvalProof<CID> pi;
    std::vector<SF> cs(coeffs, coeffs + len);
    cs.resize(n, SF::zero());
    Jac<CID> C_prime = C;
    if (w) {
        // p_bar = q * (X - z), q has d coefficients
        std::vector<SF> pbar(n, SF::zero());
        for (size_t i = 0; i < d; ++i) {
            pbar[i + 1] = pbar[i + 1] + hide_q[i];
            pbar[i] = pbar[i] - hide_q[i] * z;
        }
        Jac<CID> C_bar = pcdl_commit<CID>(pbar.data(), n, d, hide_wbar);
        tr.absorb_g(C); tr.absorb_g(C_bar);
        tr.absorb_fr(z); tr.absorb_fr(v);
        SF a = tr.challenge();
        for (size_t i = 0; i < n; ++i) cs[i] = cs[i] + pbar[i] * a;
        SF w_prime = *hide_wbar * a + *w;
        C_prime = C.add(C_bar.mul(a)).add(P.S.mul(w_prime).neg());
        pi.hiding = true; pi.C_bar = C_bar; pi.w_prime = w_prime;
    }
    tr.absorb_g(C_prime);
    tr.absorb_fr(z); tr.absorb_fr(v);
    SF xi = tr.challenge();
    Jac<CID> H_prime = P.H.mul(xi);

    std::vector<Affine<CID>> gs(P.Gs.begin(), P.Gs.begin() + n);
    std::vector<SF> zs = construct_powers(z, n);

    size_t m = n / 2;
    for (unsigned r = 0; r < lg_n; ++r) {
        SF dot_l = scalar_dot(cs.data() + m, zs.data(), m);
        Jac<CID> L = msm<CID>(gs.data(), (const u64*)(cs.data() + m), m, g_threads).add(H_prime.mul(dot_l));
        SF dot_r = scalar_dot(cs.data(), zs.data() + m, m);
        Jac<CID> R = msm<CID>(gs.data() + m, (const u64*)cs.data(), m, g_threads).add(H_prime.mul(dot_r));
        pi.Ls.push_back(L); pi.Rs.push_back(R);

        tr.absorb_fr(xi);
        tr.absorb_g(L); tr.absorb_g(R);
        SF xi_next = tr.challenge();
        if (xi_next.is_zero()) throw std::runtime_error("zero challenge");
        SF xi_inv = xi_next.inv();
        xi = xi_next;

        // fold (pcdl.rs:427-435)
        std::vector<Jac<CID>> folded(m);
        par_for_idx(m, [&](size_t j) {
            folded[j] = Jac<CID>::from_affine(gs[m + j]).mul(xi_next).add_affine(gs[j]);
        });
        std::vector<Affine<CID>> na = normalize_batch<CID>(folded);
        for (size_t j = 0; j < m; ++j) {
            gs[j] = na[j];
            cs[j] = cs[j] + cs[m + j] * xi_inv;
            zs[j] = zs[j] + zs[m + j] * xi_next;
        }
        m /= 2;
    }
    pi.U = Jac<CID>::from_affine(gs[0]);
    pi.c = cs[0];
    return pi;
}

// ---- pcdl.rs:483-554 ----  returns false on rejection
template <int CID>
bool pcdl_succinct_check(const Instance<CID>& q, HPoly<CID>* h_out, Jac<CID>* U_out, std::string* why) {
    typedef typename Jac<CID>::S SF;
    auto& P = pp<CID>();
    size_t n = q.d + 1;
    if (n & (n - 1)) throw std::runtime_error("n is not a power of two");
    unsigned lg_n = 0; while (((size_t)1 << lg_n) < n) ++lg_n;
    if (q.d > P.D()) { if (why) *why = "d was larger than D!"; return false; }
    if (q.pi.Ls.size() != lg_n || q.pi.Rs.size() != lg_n) { if (why) *why = "proof length"; return false; }
    Transcript<CID> tr(PROTO_PCDL);
    Jac<CID> C_prime = q.C;
    if (q.pi.hiding) {
        tr.absorb_g(q.C); tr.absorb_g(q.pi.C_bar);
        tr.absorb_fr(q.z); tr.absorb_fr(q.v);
        SF a = tr.challenge();
        C_prime = q.C.add(q.pi.C_bar.mul(a)).add(P.S.mul(q.pi.w_prime).neg());
    }
    tr.absorb_g(C_prime);
    tr.absorb_fr(q.z); tr.absorb_fr(q.v);
    SF xi0 = tr.challenge();
    HPoly<CID> h; h.xis.push_back(xi0);
    Jac<CID> H_prime = P.H.mul(xi0);
    Jac<CID> C_i = C_prime.add(H_prime.mul(q.v));
    for (unsigned i = 0; i < lg_n; ++i) {
        tr.absorb_fr(h.xis[i]);
        tr.absorb_g(q.pi.Ls[i]); tr.absorb_g(q.pi.Rs[i]);
        SF xi_next = tr.challenge();
        h.xis.push_back(xi_next);
        if (xi_next.is_zero()) throw std::runtime_error("zero challenge");
        C_i = C_i.add(q.pi.Ls[i].mul(xi_next.inv()).add(q.pi.Rs[i].mul(xi_next)));
    }
    SF v_prime = q.pi.c * h.eval(q.z);
    Jac<CID> rhs = q.pi.U.mul(q.pi.c).add(H_prime.mul(v_prime));
    if (!C_i.eq(rhs)) { if (why) *why = "C_(log_n) != CM.Commit(c || v')"; return false; }
    if (h_out) *h_out = h;
    if (U_out) *U_out = q.pi.U;
    return true;
}

// ---- pcdl.rs:563-583 ----
template <int CID>
bool pcdl_check(const Instance<CID>& q, std::string* why) {
    HPoly<CID> h; Jac<CID> U;
    if (!pcdl_succinct_check<CID>(q, &h, &U, why)) return false;
    auto hp = h.get_poly();
    size_t len = poly_len(hp.data(), hp.size());
    Jac<CID> comm = pedersen_commit<CID>(nullptr, pp<CID>().Gs.data(), q.d + 1, hp.data(), len);
    if (!U.eq(comm)) { if (why) *why = "U != CM.Commit(ck, h_vec)"; return false; }
    return true;
}

// ---- acc.rs:128-176 ----
template <int CID>
struct AccCommon {
    typedef typename Jac<CID>::S SF;
    Jac<CID> C_bar; size_t d; SF z;
    std::vector<HPoly<CID>> hs; std::vector<SF> alphas;
    SF eval(const SF& x) const {
        SF v = SF::zero();
        for (size_t i = 0; i < hs.size(); ++i) v = v + hs[i].eval(x) * alphas[i];
        return v;
    }
    std::vector<SF> get_poly() const {
        std::vector<SF> h;
        for (size_t i = 0; i < hs.size(); ++i) {
            auto hi = hs[i].get_poly();
            if (h.size() < hi.size()) h.resize(hi.size(), SF::zero());
            for (size_t j = 0; j < hi.size(); ++j) h[j] = h[j] + hi[j] * alphas[i];
        }
        return h;
    }
};

template <int CID>
bool acc_common_subroutine(const std::vector<Instance<CID>>& qs, AccCommon<CID>* out, std::string* why) {
    typedef typename Jac<CID>::S SF;
    if (qs.empty()) { if (why) *why = "No instances given"; return false; }
    size_t m = qs.size();
    size_t d = qs[0].d;
    Transcript<CID> tr(PROTO_ASDL);
    std::vector<Jac<CID>> Us;
    out->hs.clear();
    for (auto& q : qs) {
        HPoly<CID> h; Jac<CID> U;
        if (!pcdl_succinct_check<CID>(q, &h, &U, why)) return false;
        out->hs.push_back(h); Us.push_back(U);
        if (q.d != d) { if (why) *why = "d_i != d"; return false; }
    }
    for (auto& h : out->hs) for (auto& x : h.xis) tr.absorb_fr(x);
    for (auto& U : Us) tr.absorb_g(U);
    SF alpha = tr.challenge();
    out->alphas = construct_powers(alpha, m);
    // point_dot (group.rs:53-56): normalize_batch + MSM
    auto Ua = normalize_batch<CID>(Us);
    out->C_bar = msm<CID>(Ua.data(), (const u64*)out->alphas.data(), m, g_threads);
    out->z = tr.challenge();
    out->d = d;
    return true;
}

// ---- acc.rs:178-198 ----
template <int CID>
bool acc_prover(const std::vector<Instance<CID>>& qs, Instance<CID>* acc, std::string* why) {
    typedef typename Jac<CID>::S SF;
    AccCommon<CID> cs;
    if (!acc_common_subroutine<CID>(qs, &cs, why)) return false;
    SF v = cs.eval(cs.z);
    auto h = cs.get_poly();
    acc->C = cs.C_bar; acc->d = cs.d; acc->z = cs.z; acc->v = v;
    // pcdl::open evaluates p(z) itself (pcdl.rs:471); the instance keeps h.eval(z)
    SF v_open = poly_eval(h.data(), h.size(), cs.z);
    acc->pi = pcdl_open_without_eval<CID>(h.data(), h.size(), cs.C_bar, cs.d, cs.z, v_open, nullptr, nullptr, nullptr);
    return true;
}

// ---- acc.rs:200-213 ----
template <int CID>
bool acc_verifier(const std::vector<Instance<CID>>& qs, const Instance<CID>& acc, std::string* why) {
    AccCommon<CID> cs;
    if (!acc_common_subroutine<CID>(qs, &cs, why)) return false;
    if (!cs.C_bar.eq(acc.C)) { if (why) *why = "C_bar' != C_bar"; return false; }
    if (cs.z != acc.z) { if (why) *why = "z' = z"; return false; }
    if (cs.d != acc.d) { if (why) *why = "d' = d"; return false; }
    if (cs.eval(acc.z) != acc.v) { if (why) *why = "h(z) = v"; return false; }
    return true;
}

}  // namespace orc
