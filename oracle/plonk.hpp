// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// CPU restatement of the reference's PLONK prover and verifier for the path that drives the MSM / NTT hot
// kernels:
//   crates/plonk/src/plonk/protocol.rs:64-355    PlonkProof::naive_prover, round by round, literally: polynomial
//       products as FFT products, f_GC through ~150 pointwise passes over the 8n domain, z as a serial running
//       product with one division per row, t = f / Z_H by coefficient-form division
//   crates/plonk/src/plonk/protocol.rs:357-502   verify_succinct / verify
//   crates/plonk/src/plonk/protocol.rs:505-548   deg0, t_split, geometric_polys
//   crates/plonk/src/plonk/protocol.rs:591-1011  the five custom-gate constraint expressions
//   crates/group/src/poly.rs:21-64,133-139       Evals index convention (row i sits at omega^(i+1))
// Parity status: UNPINNED by the reference — it ships no stored proof and cannot be run here (no Rust toolchain).
// The restatement is anchored through the pinned primitives it calls (transcript, commit, open, accumulate, NTT).
#pragma once
#include "pcdl.hpp"

namespace orc {
namespace plonk {

enum { T_POLYS = 16, W_POLYS = 16, R_POLYS = 15, Q_POLYS = 10, S_POLYS = 8, DEG_MULT = 8 };
enum { N_COLUMNS = Q_POLYS + W_POLYS + R_POLYS + 2 * S_POLYS };   // flat input order: qs, ws, rs, ids, sigmas

static inline size_t evalproof_words(unsigned lg_n) { return 33 + 24 * (size_t)lg_n; }
// vs: ws rs qs ts ids sigmas z z_omega w_omegas (78 scalars) | Cs: ws ts z (33 points) | pi_r | pi_r_omega | acc_next
static inline size_t proof_words(unsigned lg_n) {
    return 78 * 4 + 33 * 12 + 2 * evalproof_words(lg_n) + instance_words(lg_n);
}

template <class SF> using Poly = std::vector<SF>;

template <class SF> void strip(Poly<SF>& p) { while (!p.empty() && p.back().is_zero()) p.pop_back(); }

template <class SF> Poly<SF> poly_add(const Poly<SF>& a, const Poly<SF>& b) {
    Poly<SF> r(std::max(a.size(), b.size()), SF::zero());
    for (size_t i = 0; i < a.size(); ++i) r[i] = a[i];
    for (size_t i = 0; i < b.size(); ++i) r[i] = r[i] + b[i];
    strip(r); return r;
}
template <class SF> Poly<SF> poly_sub(const Poly<SF>& a, const Poly<SF>& b) {
    Poly<SF> r(std::max(a.size(), b.size()), SF::zero());
    for (size_t i = 0; i < a.size(); ++i) r[i] = a[i];
    for (size_t i = 0; i < b.size(); ++i) r[i] = r[i] - b[i];
    strip(r); return r;
}
template <class SF> Poly<SF> poly_scale(const Poly<SF>& a, const SF& s) {
    Poly<SF> r(a.size());
    for (size_t i = 0; i < a.size(); ++i) r[i] = a[i] * s;
    strip(r); return r;
}
template <class SF> Poly<SF> deg0(const SF& x) { Poly<SF> r(1, x); strip(r); return r; }

// `&a * &b` on DensePolynomial: FFT product at the next power of two
template <int FID> Poly<Fe<FID>> poly_mul(const Poly<Fe<FID>>& a, const Poly<Fe<FID>>& b) {
    typedef Fe<FID> SF;
    if (a.empty() || b.empty()) return Poly<SF>();
    size_t len = a.size() + b.size() - 1;
    unsigned lg = 0; while (((size_t)1 << lg) < len) ++lg;
    size_t N = (size_t)1 << lg;
    std::vector<SF> x(a), y(b);
    x.resize(N, SF::zero()); y.resize(N, SF::zero());
    ntt_inplace<FID>(x.data(), lg, false, nullptr, g_threads);
    ntt_inplace<FID>(y.data(), lg, false, nullptr, g_threads);
    for (size_t i = 0; i < N; ++i) x[i] = x[i] * y[i];
    ntt_inplace<FID>(x.data(), lg, true, nullptr, g_threads);
    x.resize(len); strip(x); return x;
}

// evaluate_over_domain_by_ref: zero-pad, or reduce mod X^N - 1 when the polynomial is longer than the domain
template <int FID> std::vector<Fe<FID>> evaluate_over_domain(const Poly<Fe<FID>>& p, unsigned lg) {
    typedef Fe<FID> SF;
    size_t N = (size_t)1 << lg;
    std::vector<SF> v(N, SF::zero());
    for (size_t i = 0; i < p.size(); ++i) v[i & (N - 1)] = v[i & (N - 1)] + p[i];
    ntt_inplace<FID>(v.data(), lg, false, nullptr, g_threads);
    return v;
}
template <int FID> Poly<Fe<FID>> interpolate(std::vector<Fe<FID>> evals, unsigned lg) {
    ntt_inplace<FID>(evals.data(), lg, true, nullptr, g_threads);
    strip(evals); return evals;
}
template <class SF> std::vector<SF> rotate_right1(const std::vector<SF>& v) {      // Evals::from_vec_and_domain
    std::vector<SF> r(v.size());
    for (size_t i = 0; i < v.size(); ++i) r[(i + 1) % v.size()] = v[i];
    return r;
}
template <class SF> std::vector<SF> rotate_left(const std::vector<SF>& v, size_t k) {   // shift_left*(…)
    std::vector<SF> r(v.size());
    for (size_t i = 0; i < v.size(); ++i) r[i] = v[(i + k) % v.size()];
    return r;
}
template <class SF> SF eval_at(const Poly<SF>& p, const SF& x) { return poly_eval(p.data(), p.size(), x); }
template <class SF> SF pow_u64(SF b, u64 e) {
    SF r = SF::one();
    while (e) { if (e & 1) r = r * b; b = b * b; e >>= 1; }
    return r;
}

// ---- constraint expressions, written once over a ring-like type (protocol.rs:591-1011) ----
template <class T> T pow7(const T& w) { T w2 = w * w; T w3 = w2 * w; return w3 * w3 * w; }

template <class T, class SF>
T poseidon_terms(const SF* M /*3x3*/, const T* r, const T* w, const T* nw) {
    auto round = [&](const T& w0, const T& w1, const T& w2, const T& w3, const T& w4, const T& w5,
                     const T& r0, const T& r1, const T& r2) {
        T s0 = pow7(w0), s1 = pow7(w1), s2 = pow7(w2);
        return w3 - (r0 + s0 * M[0] + s1 * M[1] + s2 * M[2])
             + w4 - (r1 + s0 * M[3] + s1 * M[4] + s2 * M[5])
             + w5 - (r2 + s0 * M[6] + s1 * M[7] + s2 * M[8]);
    };
    return round(w[0], w[1], w[2], w[3], w[4], w[5], r[0], r[1], r[2])
         + round(w[3], w[4], w[5], w[6], w[7], w[8], r[3], r[4], r[5])
         + round(w[6], w[7], w[8], w[9], w[10], w[11], r[6], r[7], r[8])
         + round(w[9], w[10], w[11], w[12], w[13], w[14], r[9], r[10], r[11])
         + round(w[12], w[13], w[14], nw[0], nw[1], nw[2], r[12], r[13], r[14]);
}

// the complete affine addition (xp,yp) + (xq,yq) = (xr,yr) with witnesses alpha, beta, gamma, delta, lambda
template <class T>
T complete_add_terms(const T& one, const T& xp, const T& yp, const T& xq, const T& yq, const T& xr, const T& yr,
                     const T& al, const T& be, const T& ga, const T& de, const T& la) {
    T dx = xq - xp, dy = yq - yp;
    T result = dx * (dx * la - dy);
    T xpxp = xp * xp;
    result = result + (one - dx * al) * ((yp + yp) * la - (xpxp + xpxp + xpxp));
    T xpxq = xp * xq;
    T a = xpxq * dx;
    T b = la * la - xp - xq - xr;
    T c = la * (xp - xr) - yp - yr;
    T sy = yq + yp;
    T d = xpxq * sy;
    result = result + a * b + a * c + d * b + d * c;
    T e = one - xp * be;
    result = result + e * (xr - xq) + e * (yr - yq);
    T f = one - xq * ga;
    result = result + f * (xr - xp) + f * (yr - yp);
    T g = one - dx * al - sy * de;
    return result + g * xr + g * yr;
}

template <class T> T affine_add_terms(const T& one, const T* w) {
    return complete_add_terms(one, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], w[8], w[9], w[10]);
}

template <class T> T affine_mul_terms(const T& one, const T* w, const T* nw, const T& two_pow_i) {
    const T &xp = w[0], &yp = w[1], &a = w[2], &xg = w[3], &yg = w[4], &b = w[5], &xq = w[6], &yq = w[7], &xr = w[8],
            &yr = w[9], &bq = w[10], &lq = w[11], &ar = w[12], &gr = w[13], &dr = w[14], &lr = w[15];
    T xpxp = xp * xp;
    T e = one - xp * bq;
    T result = e * xq + e * yq;
    result = result + ((yp + yp) * lq - (xpxp + xpxp + xpxp));
    result = result + (lq * lq - (xp + xp) - xq);
    result = result + (lq * (xp - xq) - yp - yq);
    // Q + G = R: the add terms with (xq,yq),(xg,yg),(xr,yr); the two identity-case rows use 1 - xp*bq and 1 - xg*gr
    T dx = xg - xq, dy = yg - yq;
    result = result + dx * (dx * lr - dy);
    T xqxq = xq * xq;
    result = result + (one - dx * ar) * ((yq + yq) * lr - (xqxq + xqxq + xqxq));
    T xqxg = xq * xg;
    T u = xqxg * dx;
    T v = lr * lr - xq - xg - xr;
    T c = lr * (xq - xr) - yq - yr;
    T sy = yg + yq;
    T d = xqxg * sy;
    result = result + u * v + u * c + d * v + d * c;
    result = result + e * (xr - xg) + e * (yr - yg);
    T f = one - xg * gr;
    result = result + f * (xr - xq) + f * (yr - yq);
    T g = one - dx * ar - sy * dr;
    result = result + g * xr + g * yr;
    result = result + b * (b - one);
    result = result + (nw[0] - (b * xr + (one - b) * xq));
    result = result + (nw[1] - (b * yr + (one - b) * yq));
    return result + nw[2] - (a + b * two_pow_i);
}

template <class T> T range_check_terms(const T* w, const T* nw, const T* r) {
    T result = nw[0] - w[0];
    for (int i = 0; i < 15; ++i) result = result - w[1 + i] * r[i];
    return result;
}

template <class T> T eq_terms(const T* w) {
    T d = w[0] - w[1];
    return d * w[3] + (d * w[4] + w[3] - w[2]);
}

// f_GC at one point given the 10 selector, 16 witness, 15 round-constant, 3 next-row values and PI (protocol.rs:183-193)
template <class SF>
SF f_gc_point(const SF* M, const SF* q, const SF* w, const SF* r, const SF* nw, const SF& pi) {
    SF one = SF::one();
    return w[0] * q[0] + q[1] * w[1] + q[2] * w[2] + q[3] * w[0] * w[1] + q[4]
         + q[5] * poseidon_terms<SF, SF>(M, r, w, nw)
         + q[6] * affine_add_terms<SF>(one, w)
         + q[7] * affine_mul_terms<SF>(one, w, nw, r[0])
         + q[8] * eq_terms<SF>(w)
         + q[9] * range_check_terms<SF>(w, nw, r)
         + pi;
}

template <int CID>
struct Proof {
    typedef typename Jac<CID>::S SF;
    SF v_ws[W_POLYS], v_rs[R_POLYS], v_qs[Q_POLYS], v_ts[T_POLYS], v_ids[S_POLYS], v_sigmas[S_POLYS];
    SF v_z, v_z_omega, v_w_omegas[3];
    Jac<CID> C_ws[W_POLYS], C_ts[T_POLYS], C_z;
    EvalProof<CID> pi_r, pi_r_omega;
    Instance<CID> acc_next;
};

template <int CID> void evalproof_to_flat(const EvalProof<CID>& pi, unsigned L, u64* w) {
    for (unsigned i = 0; i < L; ++i) { pi.Ls[i].to_raw(w); w += 12; }
    for (unsigned i = 0; i < L; ++i) { pi.Rs[i].to_raw(w); w += 12; }
    pi.U.to_raw(w); w += 12;
    pi.c.to_raw(w); w += 4;
    *w++ = pi.hiding ? 1 : 0;
    if (pi.hiding) { pi.C_bar.to_raw(w); w += 12; pi.w_prime.to_raw(w); } else memset(w, 0, 16 * 8);
}
template <int CID> EvalProof<CID> evalproof_from_flat(const u64* w, unsigned L) {
    typedef typename Jac<CID>::S SF;
    EvalProof<CID> pi;
    pi.Ls.resize(L); pi.Rs.resize(L);
    for (unsigned i = 0; i < L; ++i) { pi.Ls[i] = Jac<CID>::from_raw(w); w += 12; }
    for (unsigned i = 0; i < L; ++i) { pi.Rs[i] = Jac<CID>::from_raw(w); w += 12; }
    pi.U = Jac<CID>::from_raw(w); w += 12;
    pi.c = SF::from_raw(w); w += 4;
    pi.hiding = *w++ != 0;
    pi.C_bar = Jac<CID>::from_raw(w); w += 12;
    pi.w_prime = SF::from_raw(w);
    return pi;
}
template <int CID> void proof_to_flat(const Proof<CID>& p, unsigned L, u64* w) {
    auto put = [&](const typename Jac<CID>::S* s, int k) { for (int i = 0; i < k; ++i) { s[i].to_raw(w); w += 4; } };
    put(p.v_ws, W_POLYS); put(p.v_rs, R_POLYS); put(p.v_qs, Q_POLYS); put(p.v_ts, T_POLYS);
    put(p.v_ids, S_POLYS); put(p.v_sigmas, S_POLYS); put(&p.v_z, 1); put(&p.v_z_omega, 1); put(p.v_w_omegas, 3);
    for (int i = 0; i < W_POLYS; ++i) { p.C_ws[i].to_raw(w); w += 12; }
    for (int i = 0; i < T_POLYS; ++i) { p.C_ts[i].to_raw(w); w += 12; }
    p.C_z.to_raw(w); w += 12;
    evalproof_to_flat<CID>(p.pi_r, L, w); w += evalproof_words(L);
    evalproof_to_flat<CID>(p.pi_r_omega, L, w); w += evalproof_words(L);
    instance_to_flat<CID>(p.acc_next, w);
}
template <int CID> Proof<CID> proof_from_flat(const u64* w, unsigned L) {
    typedef typename Jac<CID>::S SF;
    Proof<CID> p;
    auto get = [&](SF* s, int k) { for (int i = 0; i < k; ++i) { s[i] = SF::from_raw(w); w += 4; } };
    get(p.v_ws, W_POLYS); get(p.v_rs, R_POLYS); get(p.v_qs, Q_POLYS); get(p.v_ts, T_POLYS);
    get(p.v_ids, S_POLYS); get(p.v_sigmas, S_POLYS); get(&p.v_z, 1); get(&p.v_z_omega, 1); get(p.v_w_omegas, 3);
    for (int i = 0; i < W_POLYS; ++i) { p.C_ws[i] = Jac<CID>::from_raw(w); w += 12; }
    for (int i = 0; i < T_POLYS; ++i) { p.C_ts[i] = Jac<CID>::from_raw(w); w += 12; }
    p.C_z = Jac<CID>::from_raw(w); w += 12;
    p.pi_r = evalproof_from_flat<CID>(w, L); w += evalproof_words(L);
    p.pi_r_omega = evalproof_from_flat<CID>(w, L); w += evalproof_words(L);
    p.acc_next = instance_from_flat<CID>(w);
    return p;
}

// ---- inputs as the reference's (PlonkCircuit, PlonkPublicInputs, PlonkWitness) triple carries them ----
template <int CID>
struct ProverInput {
    typedef typename Jac<CID>::S SF;
    unsigned lg_n;
    Poly<SF> qs[Q_POLYS], ws[W_POLYS], rs[R_POLYS], ids[S_POLYS], sigmas[S_POLYS];   // witness.polys
    std::vector<SF> w_evals[3];          // witness.w_evals[0..3] (Evals convention: row i at omega^(i+1))
    std::vector<SF> public_inputs;
    Instance<CID> acc_prev;
};

template <int CID>
ProverInput<CID> input_from_flat(unsigned lg_n, const u64* polys, const u64* w_evals, const u64* pis, size_t n_pi,
                                 const u64* acc_prev) {
    typedef typename Jac<CID>::S SF;
    ProverInput<CID> in;
    in.lg_n = lg_n;
    size_t n = (size_t)1 << lg_n;
    const SF* c = (const SF*)polys;
    auto take = [&](Poly<SF>* dst, int k) { for (int i = 0; i < k; ++i) { dst[i].assign(c, c + n); strip(dst[i]); c += n; } };
    take(in.qs, Q_POLYS); take(in.ws, W_POLYS); take(in.rs, R_POLYS); take(in.ids, S_POLYS); take(in.sigmas, S_POLYS);
    for (int i = 0; i < 3; ++i) in.w_evals[i].assign((const SF*)w_evals + i * n, (const SF*)w_evals + (i + 1) * n);
    in.public_inputs.assign((const SF*)pis, (const SF*)pis + n_pi);
    in.acc_prev = instance_from_flat<CID>(acc_prev);
    return in;
}

// protocol.rs:64-355
template <int CID>
Proof<CID> naive_prover(const ProverInput<CID>& in) {
    typedef typename Jac<CID>::S SF;
    enum { FID = CurveTraits<CID>::SCALAR };
    const unsigned lg = in.lg_n, lg8 = lg + 3;
    const size_t n = (size_t)1 << lg, N8 = n * DEG_MULT;
    const size_t d = n - 1;
    const SF* M = (const SF*)POSEIDON_MDS[FID];
    Transcript<CID> tr(PROTO_PLONK);
    Proof<CID> out;

    // ---- round 0 (:74-109)
    std::vector<SF> pi_vec(n, SF::zero());
    if (in.public_inputs.size() > n) throw std::runtime_error("more public inputs than rows");
    for (size_t i = 0; i < in.public_inputs.size(); ++i) pi_vec[i] = in.public_inputs[i].neg();
    Poly<SF> pi_poly = interpolate<FID>(rotate_right1(pi_vec), lg);
    Poly<SF> w_omegas[3];
    for (int i = 0; i < 3; ++i) w_omegas[i] = interpolate<FID>(rotate_left(in.w_evals[i], 1), lg);
    std::vector<std::vector<SF>> q_ev(Q_POLYS), w_ev(W_POLYS), r_ev(R_POLYS), nw_ev(3);
    for (int i = 0; i < Q_POLYS; ++i) q_ev[i] = evaluate_over_domain<FID>(in.qs[i], lg8);
    for (int i = 0; i < W_POLYS; ++i) w_ev[i] = evaluate_over_domain<FID>(in.ws[i], lg8);
    for (int i = 0; i < R_POLYS; ++i) r_ev[i] = evaluate_over_domain<FID>(in.rs[i], lg8);
    for (int i = 0; i < 3; ++i) nw_ev[i] = rotate_left(w_ev[i], N8 / n);          // shift_left_small_domain
    std::vector<SF> pi_ev = evaluate_over_domain<FID>(pi_poly, lg8);

    // ---- round 1 (:114-115)
    for (int i = 0; i < W_POLYS; ++i) {
        out.C_ws[i] = pcdl_commit<CID>(in.ws[i].data(), in.ws[i].size(), d, nullptr);
        tr.absorb_g(out.C_ws[i]);
    }

    // ---- round 3 (:129-162)
    SF beta = tr.challenge(), gamma = tr.challenge();
    Poly<SF> f_prime, g_prime;
    for (int i = 0; i < S_POLYS; ++i) {
        Poly<SF> fi = poly_add(poly_add(in.ws[i], poly_scale(in.ids[i], beta)), deg0(gamma));
        Poly<SF> gi = poly_add(poly_add(in.ws[i], poly_scale(in.sigmas[i], beta)), deg0(gamma));
        f_prime = i ? poly_mul<FID>(f_prime, fi) : fi;
        g_prime = i ? poly_mul<FID>(g_prime, gi) : gi;
    }
    std::vector<SF> fp_ev = evaluate_over_domain<FID>(f_prime, lg), gp_ev = evaluate_over_domain<FID>(g_prime, lg);
    std::vector<SF> zv(n);
    for (size_t i = 0; i < n; ++i) {
        if (i == 0) zv[i] = SF::one();
        else {
            if (gp_ev[i].is_zero()) throw std::runtime_error("division by zero in the grand product");
            zv[i] = zv[i - 1] * (fp_ev[i] * gp_ev[i].inv());
        }
    }
    std::vector<SF> z_evals = rotate_right1(zv);
    Poly<SF> z_omega = interpolate<FID>(rotate_left(z_evals, 1), lg);
    Poly<SF> z = interpolate<FID>(z_evals, lg);
    out.C_z = pcdl_commit<CID>(z.data(), z.size(), d, nullptr);
    tr.absorb_g(out.C_z);

    // ---- round 4 (:170-268)
    SF alpha = tr.challenge();
    std::vector<SF> f_gc_ev(N8);
    par_for_idx(N8, [&](size_t j) {
        SF q[Q_POLYS], w[W_POLYS], r[R_POLYS], nw[3];
        for (int i = 0; i < Q_POLYS; ++i) q[i] = q_ev[i][j];
        for (int i = 0; i < W_POLYS; ++i) w[i] = w_ev[i][j];
        for (int i = 0; i < R_POLYS; ++i) r[i] = r_ev[i][j];
        for (int i = 0; i < 3; ++i) nw[i] = nw_ev[i][j];
        f_gc_ev[j] = f_gc_point<SF>(M, q, w, r, nw, pi_ev[j]);
    });
    Poly<SF> f_gc = interpolate<FID>(f_gc_ev, lg8);
    std::vector<SF> l1_ev(n, SF::zero()); l1_ev[0] = SF::one();
    Poly<SF> l1 = interpolate<FID>(rotate_right1(l1_ev), lg);
    Poly<SF> f_cc1 = poly_mul<FID>(l1, poly_sub(z, deg0(SF::one())));
    Poly<SF> f_cc2 = poly_sub(poly_mul<FID>(z, f_prime), poly_mul<FID>(z_omega, g_prime));
    Poly<SF> f = poly_add(poly_add(f_gc, poly_scale(f_cc1, alpha)), poly_scale(f_cc2, alpha * alpha));
    // divide_by_vanishing_poly: quotient by X^n - 1, remainder dropped
    Poly<SF> t;
    if (f.size() >= n) {
        t.assign(f.begin() + n, f.end());
        for (size_t i = 1; i < f.size() / n; ++i)
            for (size_t j = 0; n * (i + 1) + j < f.size(); ++j) t[j] = t[j] + f[n * (i + 1) + j];
        strip(t);
    }
    if (t.size() > (size_t)T_POLYS * n) throw std::runtime_error("t.degree() < T_POLYS * n");
    Poly<SF> ts[T_POLYS];
    {
        std::vector<SF> tt(t); tt.resize((size_t)T_POLYS * n, SF::zero());
        for (int i = 0; i < T_POLYS; ++i) { ts[i].assign(tt.begin() + i * n, tt.begin() + (i + 1) * n); strip(ts[i]); }
    }
    for (int i = 0; i < T_POLYS; ++i) {
        out.C_ts[i] = pcdl_commit<CID>(ts[i].data(), ts[i].size(), d, nullptr);
        tr.absorb_g(out.C_ts[i]);
    }

    // ---- round 5 (:273-333)
    SF zeta = tr.challenge();
    auto geometric = [&](const std::vector<const Poly<SF>*>& ps) {
        Poly<SF> r; SF c = SF::one();
        for (auto* p : ps) { r = poly_add(r, poly_scale(*p, c)); c = c * zeta; }
        return r;
    };
    std::vector<const Poly<SF>*> v1, v2;
    for (int i = 0; i < Q_POLYS; ++i) v1.push_back(&in.qs[i]);
    for (int i = 0; i < W_POLYS; ++i) v1.push_back(&in.ws[i]);
    for (int i = 0; i < T_POLYS; ++i) v1.push_back(&ts[i]);
    v1.push_back(&z);
    for (int i = 0; i < 3; ++i) v2.push_back(&in.ws[i]);
    v2.push_back(&z);
    Poly<SF> r = geometric(v1), r_omega = geometric(v2);
    SF xi = tr.challenge();
    SF omega = SF::from_raw(root_of_unity<FID>(lg).l);
    SF xi_omega = xi * omega;
    auto open = [&](const Poly<SF>& p, const SF& at) {
        Instance<CID> q;
        q.C = pcdl_commit<CID>(p.data(), p.size(), d, nullptr);
        q.d = d; q.z = at; q.v = eval_at(p, at);
        q.pi = pcdl_open_without_eval<CID>(p.data(), p.size(), q.C, d, at, q.v, nullptr, nullptr, nullptr);
        return q;
    };
    Instance<CID> q_r = open(r, xi), q_ro = open(r_omega, xi_omega);
    std::string why;
    if (!acc_prover<CID>({in.acc_prev, q_r, q_ro}, &out.acc_next, &why)) throw std::runtime_error("acc::prover: " + why);
    for (int i = 0; i < W_POLYS; ++i) out.v_ws[i] = eval_at(in.ws[i], xi);
    for (int i = 0; i < R_POLYS; ++i) out.v_rs[i] = eval_at(in.rs[i], xi);
    for (int i = 0; i < Q_POLYS; ++i) out.v_qs[i] = eval_at(in.qs[i], xi);
    for (int i = 0; i < T_POLYS; ++i) out.v_ts[i] = eval_at(ts[i], xi);
    for (int i = 0; i < S_POLYS; ++i) { out.v_ids[i] = eval_at(in.ids[i], xi); out.v_sigmas[i] = eval_at(in.sigmas[i], xi); }
    out.v_z = eval_at(z, xi);
    out.v_z_omega = eval_at(z, xi_omega);
    for (int i = 0; i < 3; ++i) out.v_w_omegas[i] = eval_at(w_omegas[i], xi);
    out.pi_r = q_r.pi; out.pi_r_omega = q_ro.pi;
    return out;
}

// protocol.rs:357-491.  C_qs = circuit.Cs.qs; returns false with a reason on rejection.
template <int CID>
bool verify_succinct(const Proof<CID>& pi, unsigned lg, const Jac<CID>* C_qs, const typename Jac<CID>::S* pub, size_t n_pub,
                     size_t public_input_count, const Instance<CID>& acc_prev, std::string* why) {
    typedef typename Jac<CID>::S SF;
    enum { FID = CurveTraits<CID>::SCALAR };
    const size_t rows = (size_t)1 << lg, d = rows - 1;
    const SF nS = SF::from_u64(rows), one = SF::one();
    const SF* M = (const SF*)POSEIDON_MDS[FID];
    Transcript<CID> tr(PROTO_PLONK);
    if (n_pub != public_input_count) { if (why) *why = "public input count"; return false; }
    for (int i = 0; i < W_POLYS; ++i) tr.absorb_g(pi.C_ws[i]);
    SF beta = tr.challenge(), gamma = tr.challenge();
    tr.absorb_g(pi.C_z);
    SF alpha = tr.challenge();
    for (int i = 0; i < T_POLYS; ++i) tr.absorb_g(pi.C_ts[i]);
    SF zeta = tr.challenge(), xi = tr.challenge();
    SF xi_n = xi; for (unsigned i = 0; i < lg; ++i) xi_n = xi_n * xi_n;
    SF omega = SF::from_raw(root_of_unity<FID>(lg).l);
    SF xi_omega = xi * omega;
    SF f_prime = pi.v_ws[0] + beta * pi.v_ids[0] + gamma, g_prime = pi.v_ws[0] + beta * pi.v_sigmas[0] + gamma;
    for (int i = 1; i < S_POLYS; ++i) {
        f_prime = f_prime * (pi.v_ws[i] + beta * pi.v_ids[i] + gamma);
        g_prime = g_prime * (pi.v_ws[i] + beta * pi.v_sigmas[i] + gamma);
    }
    // public_input_eval_generic (:565-589)
    SF pi_xi = SF::zero(), omega_j = omega;
    for (size_t j = 0; j < n_pub; ++j) {
        SF den = nS * (xi - omega_j);
        if (den.is_zero()) throw std::runtime_error("division by zero in the public-input evaluation");
        pi_xi = pi_xi + ((xi_n - one) * omega_j) * den.inv() * pub[j].neg();
        omega_j = omega_j * omega;
    }
    SF f_gc = f_gc_point<SF>(M, pi.v_qs, pi.v_ws, pi.v_rs, pi.v_w_omegas, pi_xi);
    SF den = nS * (xi - omega);
    if (den.is_zero()) throw std::runtime_error("division by zero in l1");
    SF l1 = (omega * (xi_n - one)) * den.inv();
    SF z_H = xi_n - one;
    SF f = f_gc + alpha * (l1 * (pi.v_z - one)) + alpha * alpha * (pi.v_z * f_prime - pi.v_z_omega * g_prime);
    SF t = pi.v_ts[0], acc = xi_n;
    for (int i = 1; i < T_POLYS; ++i) { t = t + acc * pi.v_ts[i]; acc = acc * xi_n; }
    if (f != t * z_H) { if (why) *why = "T(xi) != F(xi) / Z_H(xi)"; return false; }

    auto geo_s = [&](const std::vector<SF>& v) { SF r = v[0], a = zeta; for (size_t i = 1; i < v.size(); ++i) { r = r + v[i] * a; a = a * zeta; } return r; };
    auto geo_p = [&](const std::vector<Jac<CID>>& v) { Jac<CID> r = v[0]; SF a = zeta; for (size_t i = 1; i < v.size(); ++i) { r = r.add(v[i].mul(a)); a = a * zeta; } return r; };
    std::vector<SF> sv; std::vector<Jac<CID>> pv;
    for (int i = 0; i < Q_POLYS; ++i) { sv.push_back(pi.v_qs[i]); pv.push_back(C_qs[i]); }
    for (int i = 0; i < W_POLYS; ++i) { sv.push_back(pi.v_ws[i]); pv.push_back(pi.C_ws[i]); }
    for (int i = 0; i < T_POLYS; ++i) { sv.push_back(pi.v_ts[i]); pv.push_back(pi.C_ts[i]); }
    sv.push_back(pi.v_z); pv.push_back(pi.C_z);
    SF v_r = geo_s(sv); Jac<CID> C_r = geo_p(pv);
    SF v_ro = geo_s({pi.v_w_omegas[0], pi.v_w_omegas[1], pi.v_w_omegas[2], pi.v_z_omega});
    Jac<CID> C_ro = geo_p({pi.C_ws[0], pi.C_ws[1], pi.C_ws[2], pi.C_z});
    Instance<CID> i1, i2;
    i1.C = C_r; i1.d = d; i1.z = xi; i1.v = v_r; i1.pi = pi.pi_r;
    i2.C = C_ro; i2.d = d; i2.z = xi_omega; i2.v = v_ro; i2.pi = pi.pi_r_omega;
    return acc_verifier<CID>({acc_prev, i1, i2}, pi.acc_next, why);
}

template <int CID>
bool verify(const Proof<CID>& pi, unsigned lg, const Jac<CID>* C_qs, const typename Jac<CID>::S* pub, size_t n_pub,
            size_t public_input_count, const Instance<CID>& acc_prev, std::string* why) {
    if (!verify_succinct<CID>(pi, lg, C_qs, pub, n_pub, public_input_count, acc_prev, why)) return false;
    return pcdl_check<CID>(pi.acc_next, why);
}

}  // namespace plonk
}  // namespace orc
