// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Flat C entry points over the CPU restatement so that tests/ (ctypes), smoke()
// and bench.py's cpu_baseline / --impl reference legs can drive it.  Nothing in
// halo_b200/ may load this library.
//
// Public-parameter generation restates crates/group/src/main.rs:55-68,104-117:
//   gen(i) = G * (SHA3-256(le64(i) || "To understand recursion, ...") mod r),
//   S = gen(0), H = gen(1), Gs[16384*blk + k] = gen(blk + k + 2)
// and the blob reader restates the bincode `standard()` varint layout consumed
// at crates/group/src/pp.rs:36-53.
#include "pcdl.hpp"
#include "plonk.hpp"
#include "serialize.hpp"
#include <cstdio>
#include <map>

namespace orc {
unsigned g_threads = 1;

// ---------------- SHA3-256 (FIPS 202) ----------------
static inline u64 rotl64(u64 x, int s) { return s ? (x << s) | (x >> (64 - s)) : x; }
static void keccak_f(u64 st[25]) {
    static const u64 RC[24] = {
        0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
        0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
        0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
        0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
        0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
        0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
    static const int ROT[25] = {0, 1, 62, 28, 27, 36, 44, 6, 55, 20, 3, 10, 43, 25, 39, 41, 45, 15, 21, 8, 18, 2, 61, 56, 14};
    for (int round = 0; round < 24; ++round) {
        u64 C[5], D[5], B[25];
        for (int x = 0; x < 5; ++x) C[x] = st[x] ^ st[x + 5] ^ st[x + 10] ^ st[x + 15] ^ st[x + 20];
        for (int x = 0; x < 5; ++x) D[x] = C[(x + 4) % 5] ^ rotl64(C[(x + 1) % 5], 1);
        for (int i = 0; i < 25; ++i) st[i] ^= D[i % 5];
        for (int x = 0; x < 5; ++x)
            for (int y = 0; y < 5; ++y)
                B[y + 5 * ((2 * x + 3 * y) % 5)] = rotl64(st[x + 5 * y], ROT[x + 5 * y]);
        for (int y = 0; y < 5; ++y)
            for (int x = 0; x < 5; ++x)
                st[x + 5 * y] = B[x + 5 * y] ^ (~B[(x + 1) % 5 + 5 * y] & B[(x + 2) % 5 + 5 * y]);
        st[0] ^= RC[round];
    }
}
static void sha3_256(const unsigned char* msg, size_t len, unsigned char out[32]) {
    const size_t rate = 136;
    u64 st[25]; memset(st, 0, sizeof st);
    unsigned char block[136];
    while (len >= rate) {
        for (size_t i = 0; i < rate / 8; ++i) { u64 w; memcpy(&w, msg + 8 * i, 8); st[i] ^= w; }
        keccak_f(st); msg += rate; len -= rate;
    }
    memset(block, 0, rate);
    memcpy(block, msg, len);
    block[len] ^= 0x06;
    block[rate - 1] ^= 0x80;
    for (size_t i = 0; i < rate / 8; ++i) { u64 w; memcpy(&w, block + 8 * i, 8); st[i] ^= w; }
    keccak_f(st);
    memcpy(out, st, 32);
}

// from_le_bytes_mod_order for a 32-byte digest: value < 2^256 < 4p, subtract p while >= p
template <int FID>
static Fe<FID> scalar_from_digest(const unsigned char d[32]) {
    u64 w[4]; memcpy(w, d, 32);
    const u64* p = field_params(FID).p;
    while (geq4(w, p)) sub4(w, w, p);
    return Fe<FID>::from_canonical(w);
}

template <int CID>
static Jac<CID> generator_hash(u64 i) {
    static const char genesis[] = "To understand recursion, one must first understand recursion";
    unsigned char buf[8 + sizeof(genesis) - 1];
    memcpy(buf, &i, 8);
    memcpy(buf + 8, genesis, sizeof(genesis) - 1);
    unsigned char dig[32];
    sha3_256(buf, sizeof buf, dig);
    return Jac<CID>::generator().mul(scalar_from_digest<CurveTraits<CID>::SCALAR>(dig));
}

template <int CID>
static void pp_generate(size_t n, u64* gs_xy, u64* s_xy, u64* h_xy) {
    const size_t BLK = 16384;
    size_t nblocks = (n + BLK - 1) / BLK;
    size_t distinct = std::min(n, BLK) + (nblocks > 0 ? nblocks - 1 : 0);
    std::vector<Jac<CID>> pts(distinct);
    par_for_idx(distinct, [&](size_t k) { pts[k] = generator_hash<CID>(k + 2); });
    auto aff = normalize_batch<CID>(pts);
    for (size_t idx = 0; idx < n; ++idx) {
        size_t blk = idx / BLK, k = idx % BLK;
        aff[blk + k].to_raw(gs_xy + 8 * idx);
    }
    if (s_xy) generator_hash<CID>(0).to_affine().to_raw(s_xy);
    if (h_xy) generator_hash<CID>(1).to_affine().to_raw(h_xy);
}

template <int CID>
static void pp_set(const u64* gs_xy, size_t n, const u64* s_xy, const u64* h_xy) {
    auto& P = pp<CID>();
    P.Gs.resize(n);
    for (size_t i = 0; i < n; ++i) P.Gs[i] = Affine<CID>::from_raw(gs_xy + 8 * i);
    P.S = Jac<CID>::from_affine(Affine<CID>::from_raw(s_xy));
    P.H = Jac<CID>::from_affine(Affine<CID>::from_raw(h_xy));
    P.loaded = true;
}

static thread_local std::string g_err;

template <int CID>
static std::vector<Instance<CID>> load_instances(const u64* flat, size_t m, unsigned lg_n) {
    std::vector<Instance<CID>> qs;
    for (size_t i = 0; i < m; ++i) qs.push_back(instance_from_flat<CID>(flat + i * instance_words(lg_n)));
    return qs;
}

}  // namespace orc

using namespace orc;

// deterministic distinct test points: P_i = P_0 + i*D; k0, kd canonical 64-bit multipliers of G
template <int CID>
static void points_arith(u64 k0, u64 kd, size_t n, u64* out_xy) {
    u64 w0[4] = {k0, 0, 0, 0}, wd[4] = {kd, 0, 0, 0};
    Jac<CID> G = Jac<CID>::generator();
    Jac<CID> base = G.mul_bits(w0);
    Affine<CID> D = G.mul_bits(wd).to_affine();
    // chunked so that normalisation memory stays bounded; each thread walks its own range
    unsigned T = g_threads;
    std::vector<Jac<CID>> pts(n);
    par_for(n, T, [&](size_t lo, size_t hi) {
        u64 wl[4] = {(u64)lo, 0, 0, 0};
        // start = base + lo * D
        Jac<CID> cur = base.add(Jac<CID>::from_affine(D).mul_bits(wl));
        for (size_t i = lo; i < hi; ++i) { pts[i] = cur; cur = cur.add_affine(D); }
    });
    const size_t CH = 1 << 14;
    par_for((n + CH - 1) / CH, T, [&](size_t lo, size_t hi) {
        for (size_t c = lo; c < hi; ++c) {
            size_t a = c * CH, b = std::min(n, a + CH);
            std::vector<Jac<CID>> sub(pts.begin() + a, pts.begin() + b);
            auto aff = normalize_batch<CID>(sub);
            for (size_t i = a; i < b; ++i) aff[i - a].to_raw(out_xy + 8 * i);
        }
    });
}

template <int CID>
static int open_impl(const u64* coeffs, size_t len, size_t d, const u64* z, const u64* w, const u64* hq, const u64* hw, u64* out) {
    typedef typename Jac<CID>::S SF;
    SF zz = SF::from_raw(z);
    SF ww, wb; if (w) { ww = SF::from_raw(w); wb = SF::from_raw(hw); }
    Instance<CID> q;
    // Instance::open (pcdl.rs:41-52): C = commit(p,d,w); v = p(z); pi = open(..)
    q.C = pcdl_commit<CID>((const SF*)coeffs, len, d, w ? &ww : nullptr);
    q.d = d; q.z = zz;
    size_t l = poly_len((const SF*)coeffs, len);
    q.v = poly_eval((const SF*)coeffs, l, zz);
    q.pi = pcdl_open_without_eval<CID>((const SF*)coeffs, len, q.C, d, zz, q.v, w ? &ww : nullptr, (const SF*)hq, w ? &wb : nullptr);
    instance_to_flat<CID>(q, out);
    return 0;
}


template <int CID>
static int plonk_prove_impl(unsigned lg_n, const u64* polys, const u64* w_evals, const u64* pub, size_t n_pub,
                            const u64* acc_prev, u64* out) {
    auto in = plonk::input_from_flat<CID>(lg_n, polys, w_evals, pub, n_pub, acc_prev);
    auto pi = plonk::naive_prover<CID>(in);
    plonk::proof_to_flat<CID>(pi, lg_n, out);
    return 0;
}
template <int CID>
static int plonk_verify_impl(unsigned lg_n, const u64* C_qs, const u64* pub, size_t n_pub, size_t count,
                             const u64* acc_prev, const u64* proof, int succinct_only) {
    typedef typename Jac<CID>::S SF;
    auto pi = plonk::proof_from_flat<CID>(proof, lg_n);
    Jac<CID> cq[plonk::Q_POLYS];
    for (int i = 0; i < plonk::Q_POLYS; ++i) cq[i] = Jac<CID>::from_raw(C_qs + 12 * i);
    std::string why;
    auto acc = instance_from_flat<CID>(acc_prev);
    bool ok = succinct_only ? plonk::verify_succinct<CID>(pi, lg_n, cq, (const SF*)pub, n_pub, count, acc, &why)
                            : plonk::verify<CID>(pi, lg_n, cq, (const SF*)pub, n_pub, count, acc, &why);
    if (!ok) { g_err = why; return 1; }
    return 0;
}

#define DISPATCH_CURVE(curve, CALL0, CALL1) do { if ((curve) == 0) { CALL0; } else { CALL1; } } while (0)

extern "C" {

void orc_set_threads(unsigned n) { g_threads = n ? n : 1; }
const char* orc_last_error() { return g_err.c_str(); }

// ---- field ----
void orc_field_modulus(int field, u64* out) { memcpy(out, field_params(field).p, 32); }
void orc_field_mul(int field, const u64* a, const u64* b, u64* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        if (field == 0) (Fp::from_raw(a + 4 * i) * Fp::from_raw(b + 4 * i)).to_raw(out + 4 * i);
        else (Fq::from_raw(a + 4 * i) * Fq::from_raw(b + 4 * i)).to_raw(out + 4 * i);
    }
}
void orc_field_add(int field, const u64* a, const u64* b, u64* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        if (field == 0) (Fp::from_raw(a + 4 * i) + Fp::from_raw(b + 4 * i)).to_raw(out + 4 * i);
        else (Fq::from_raw(a + 4 * i) + Fq::from_raw(b + 4 * i)).to_raw(out + 4 * i);
    }
}
void orc_field_sub(int field, const u64* a, const u64* b, u64* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        if (field == 0) (Fp::from_raw(a + 4 * i) - Fp::from_raw(b + 4 * i)).to_raw(out + 4 * i);
        else (Fq::from_raw(a + 4 * i) - Fq::from_raw(b + 4 * i)).to_raw(out + 4 * i);
    }
}
void orc_field_inv(int field, const u64* a, u64* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        if (field == 0) Fp::from_raw(a + 4 * i).inv().to_raw(out + 4 * i);
        else Fq::from_raw(a + 4 * i).inv().to_raw(out + 4 * i);
    }
}
void orc_field_to_canonical(int field, const u64* a, u64* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        if (field == 0) Fp::from_raw(a + 4 * i).to_canonical(out + 4 * i);
        else Fq::from_raw(a + 4 * i).to_canonical(out + 4 * i);
    }
}
void orc_field_from_canonical(int field, const u64* a, u64* out, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        if (field == 0) Fp::from_canonical(a + 4 * i).to_raw(out + 4 * i);
        else Fq::from_canonical(a + 4 * i).to_raw(out + 4 * i);
    }
}
void orc_root_of_unity(int field, unsigned log_n, u64* out) {
    if (field == 0) root_of_unity<0>(log_n).to_raw(out); else root_of_unity<1>(log_n).to_raw(out);
}

// ---- curve ----
void orc_point_to_affine(int curve, const u64* xyz, u64* xy, size_t n) {
    for (size_t i = 0; i < n; ++i)
        DISPATCH_CURVE(curve, Jac<0>::from_raw(xyz + 12 * i).to_affine().to_raw(xy + 8 * i),
                       Jac<1>::from_raw(xyz + 12 * i).to_affine().to_raw(xy + 8 * i));
}
int orc_point_on_curve(int curve, const u64* xy) {
    int r = 0;
    DISPATCH_CURVE(curve, r = Affine<0>::from_raw(xy).on_curve(), r = Affine<1>::from_raw(xy).on_curve());
    return r;
}
void orc_point_add(int curve, const u64* a_xyz, const u64* b_xyz, u64* out_xyz) {
    DISPATCH_CURVE(curve, Jac<0>::from_raw(a_xyz).add(Jac<0>::from_raw(b_xyz)).to_raw(out_xyz),
                   Jac<1>::from_raw(a_xyz).add(Jac<1>::from_raw(b_xyz)).to_raw(out_xyz));
}
void orc_point_mul(int curve, const u64* a_xyz, const u64* scalar, u64* out_xyz) {
    DISPATCH_CURVE(curve, Jac<0>::from_raw(a_xyz).mul(Fp::from_raw(scalar)).to_raw(out_xyz),
                   Jac<1>::from_raw(a_xyz).mul(Fq::from_raw(scalar)).to_raw(out_xyz));
}
void orc_generator(int curve, u64* out_xyz) {
    DISPATCH_CURVE(curve, Jac<0>::generator().to_raw(out_xyz), Jac<1>::generator().to_raw(out_xyz));
}
void orc_points_arith(int curve, u64 k0, u64 kd, size_t n, u64* out_xy) {
    DISPATCH_CURVE(curve, points_arith<0>(k0, kd, n, out_xy), points_arith<1>(k0, kd, n, out_xy));
}

// ---- MSM ----
void orc_msm(int curve, const u64* bases_xy, const u64* scalars, size_t k, u64* out_xyz) {
    if (curve == 0) {
        std::vector<Affine<0>> b(k); for (size_t i = 0; i < k; ++i) b[i] = Affine<0>::from_raw(bases_xy + 8 * i);
        msm<0>(b.data(), scalars, k, g_threads).to_raw(out_xyz);
    } else {
        std::vector<Affine<1>> b(k); for (size_t i = 0; i < k; ++i) b[i] = Affine<1>::from_raw(bases_xy + 8 * i);
        msm<1>(b.data(), scalars, k, g_threads).to_raw(out_xyz);
    }
}
void orc_msm_naive(int curve, const u64* bases_xy, const u64* scalars, size_t k, u64* out_xyz) {
    if (curve == 0) {
        std::vector<Affine<0>> b(k); for (size_t i = 0; i < k; ++i) b[i] = Affine<0>::from_raw(bases_xy + 8 * i);
        msm_naive<0>(b.data(), scalars, k).to_raw(out_xyz);
    } else {
        std::vector<Affine<1>> b(k); for (size_t i = 0; i < k; ++i) b[i] = Affine<1>::from_raw(bases_xy + 8 * i);
        msm_naive<1>(b.data(), scalars, k).to_raw(out_xyz);
    }
}
// MSM against the loaded generators Gs[off .. off+k)
int orc_msm_pp(int curve, size_t off, const u64* scalars, size_t k, u64* out_xyz) {
    if (curve == 0) { if (off + k > pp<0>().Gs.size()) return -1; msm<0>(pp<0>().Gs.data() + off, scalars, k, g_threads).to_raw(out_xyz); }
    else { if (off + k > pp<1>().Gs.size()) return -1; msm<1>(pp<1>().Gs.data() + off, scalars, k, g_threads).to_raw(out_xyz); }
    return 0;
}

// ---- NTT ----
void orc_ntt(int field, u64* data, unsigned log_n, size_t batch, int inverse, const u64* coset) {
    size_t n = (size_t)1 << log_n;
    for (size_t b = 0; b < batch; ++b) {
        if (field == 0) { Fp g; if (coset) g = Fp::from_raw(coset); ntt_inplace<0>((Fp*)(data + 4 * n * b), log_n, inverse != 0, coset ? &g : nullptr, g_threads); }
        else { Fq g; if (coset) g = Fq::from_raw(coset); ntt_inplace<1>((Fq*)(data + 4 * n * b), log_n, inverse != 0, coset ? &g : nullptr, g_threads); }
    }
}
void orc_dft_naive(int field, const u64* in, u64* out, unsigned log_n) {
    if (field == 0) dft_naive<0>((const Fp*)in, (Fp*)out, log_n); else dft_naive<1>((const Fq*)in, (Fq*)out, log_n);
}

// ---- Poseidon ----
// plain sponge hash over `field`: absorb all inputs, squeeze once (inner_sponge.rs:315-322)
void orc_poseidon_hash(int field, const u64* inputs, size_t n, u64* out) {
    if (field == 0) { PoseidonSponge<0> s; for (size_t i = 0; i < n; ++i) s.absorb(Fp::from_raw(inputs + 4 * i)); s.squeeze().to_raw(out); }
    else { PoseidonSponge<1> s; for (size_t i = 0; i < n; ++i) s.absorb(Fq::from_raw(inputs + 4 * i)); s.squeeze().to_raw(out); }
}

// ---- public parameters ----
void orc_pp_generate(int curve, size_t n, u64* gs_xy, u64* s_xy, u64* h_xy) {
    DISPATCH_CURVE(curve, pp_generate<0>(n, gs_xy, s_xy, h_xy), pp_generate<1>(n, gs_xy, s_xy, h_xy));
}
void orc_pp_set(int curve, const u64* gs_xy, size_t n, const u64* s_xy, const u64* h_xy) {
    DISPATCH_CURVE(curve, pp_set<0>(gs_xy, n, s_xy, h_xy), pp_set<1>(gs_xy, n, s_xy, h_xy));
}
// bincode standard() varint stream of u64 -> raw words; returns words decoded or -1
long orc_bincode_decode_u64s(const unsigned char* bytes, size_t len, u64* out, size_t max_words, int has_len_prefix) {
    size_t pos = 0, nw = 0;
    auto rd = [&](u64& v) -> bool {
        if (pos >= len) return false;
        unsigned char t = bytes[pos++];
        size_t w = 0;
        if (t < 251) { v = t; return true; }
        if (t == 251) w = 2; else if (t == 252) w = 4; else if (t == 253) w = 8; else return false;
        if (pos + w > len) return false;
        v = 0; memcpy(&v, bytes + pos, w); pos += w; return true;
    };
    u64 count = 0;
    if (has_len_prefix) { if (!rd(count)) return -1; count *= 8; } else count = max_words;
    for (u64 i = 0; i < count && nw < max_words; ++i) { u64 v; if (!rd(v)) return -1; out[nw++] = v; }
    return (long)nw;
}

// ---- PCDL / ASDL ----
size_t orc_instance_words(unsigned lg_n) { return instance_words(lg_n); }

int orc_pcdl_commit(int curve, const u64* coeffs, size_t len, size_t d, const u64* w_or_null, u64* out_xyz) {
    try {
        if (curve == 0) { Fp w; if (w_or_null) w = Fp::from_raw(w_or_null); pcdl_commit<0>((const Fp*)coeffs, len, d, w_or_null ? &w : nullptr).to_raw(out_xyz); }
        else { Fq w; if (w_or_null) w = Fq::from_raw(w_or_null); pcdl_commit<1>((const Fq*)coeffs, len, d, w_or_null ? &w : nullptr).to_raw(out_xyz); }
    } catch (std::exception& e) { g_err = e.what(); return -1; }
    return 0;
}

// Instance::open.  hiding: w != NULL, then hide_q (d coefficients) and hide_wbar are the prover's randomness.
int orc_instance_open(int curve, const u64* coeffs, size_t len, size_t d, const u64* z, const u64* w,
                      const u64* hide_q, const u64* hide_wbar, u64* out_instance) {
    try { return curve == 0 ? open_impl<0>(coeffs, len, d, z, w, hide_q, hide_wbar, out_instance)
                            : open_impl<1>(coeffs, len, d, z, w, hide_q, hide_wbar, out_instance); }
    catch (std::exception& e) { g_err = e.what(); return -1; }
}

// 0 accept, 1 reject, -1 precondition failure.  xis_out: (lg_n+1) x 4 words or NULL.
int orc_pcdl_succinct_check(int curve, const u64* inst, unsigned lg_n, u64* xis_out, u64* U_out) {
    (void)lg_n;
    try {
        bool ok; std::string why;
        if (curve == 0) { HPoly<0> h; Jac<0> U; ok = pcdl_succinct_check<0>(instance_from_flat<0>(inst), &h, &U, &why);
            if (ok) { if (xis_out) memcpy(xis_out, h.xis.data(), h.xis.size() * 32); if (U_out) U.to_raw(U_out); } }
        else { HPoly<1> h; Jac<1> U; ok = pcdl_succinct_check<1>(instance_from_flat<1>(inst), &h, &U, &why);
            if (ok) { if (xis_out) memcpy(xis_out, h.xis.data(), h.xis.size() * 32); if (U_out) U.to_raw(U_out); } }
        if (!ok) { g_err = why; return 1; }
        return 0;
    } catch (std::exception& e) { g_err = e.what(); return -1; }
}
int orc_pcdl_check(int curve, const u64* inst) {
    try {
        bool ok; std::string why;
        if (curve == 0) ok = pcdl_check<0>(instance_from_flat<0>(inst), &why); else ok = pcdl_check<1>(instance_from_flat<1>(inst), &why);
        if (!ok) { g_err = why; return 1; }
        return 0;
    } catch (std::exception& e) { g_err = e.what(); return -1; }
}
void orc_hpoly_coeffs(int field_of_scalars, const u64* xis, unsigned lg_n, u64* out) {
    if (field_of_scalars == 0) { HPoly<0> h; h.xis.assign((const Fp*)xis, (const Fp*)xis + lg_n + 1); auto c = h.get_poly(); memcpy(out, c.data(), c.size() * 32); }
    else { HPoly<1> h; h.xis.assign((const Fq*)xis, (const Fq*)xis + lg_n + 1); auto c = h.get_poly(); memcpy(out, c.data(), c.size() * 32); }
}
void orc_hpoly_eval(int field_of_scalars, const u64* xis, unsigned lg_n, const u64* z, u64* out) {
    if (field_of_scalars == 0) { HPoly<0> h; h.xis.assign((const Fp*)xis, (const Fp*)xis + lg_n + 1); h.eval(Fp::from_raw(z)).to_raw(out); }
    else { HPoly<1> h; h.xis.assign((const Fq*)xis, (const Fq*)xis + lg_n + 1); h.eval(Fq::from_raw(z)).to_raw(out); }
}

int orc_acc_prover(int curve, const u64* insts, size_t m, unsigned lg_n, u64* out_acc) {
    try {
        bool ok; std::string why;
        if (curve == 0) { Instance<0> a; ok = acc_prover<0>(load_instances<0>(insts, m, lg_n), &a, &why); if (ok) instance_to_flat<0>(a, out_acc); }
        else { Instance<1> a; ok = acc_prover<1>(load_instances<1>(insts, m, lg_n), &a, &why); if (ok) instance_to_flat<1>(a, out_acc); }
        if (!ok) { g_err = why; return 1; }
        return 0;
    } catch (std::exception& e) { g_err = e.what(); return -1; }
}
int orc_acc_verifier(int curve, const u64* insts, size_t m, unsigned lg_n, const u64* acc) {
    try {
        bool ok; std::string why;
        if (curve == 0) ok = acc_verifier<0>(load_instances<0>(insts, m, lg_n), instance_from_flat<0>(acc), &why);
        else ok = acc_verifier<1>(load_instances<1>(insts, m, lg_n), instance_from_flat<1>(acc), &why);
        if (!ok) { g_err = why; return 1; }
        return 0;
    } catch (std::exception& e) { g_err = e.what(); return -1; }
}
// ---- PLONK prover / verifier (plonk.hpp) ----
size_t orc_plonk_proof_words(unsigned lg_n) { return plonk::proof_words(lg_n); }

// polys: 57 x n x 4 words (qs, ws, rs, ids, sigmas coefficient vectors, zero-padded); w_evals: 3 x n x 4
int orc_plonk_prove(int curve, unsigned lg_n, const u64* polys, const u64* w_evals, const u64* pub, size_t n_pub,
                    const u64* acc_prev, u64* out_proof) {
    try { return curve == 0 ? plonk_prove_impl<0>(lg_n, polys, w_evals, pub, n_pub, acc_prev, out_proof)
                            : plonk_prove_impl<1>(lg_n, polys, w_evals, pub, n_pub, acc_prev, out_proof); }
    catch (std::exception& e) { g_err = e.what(); return -1; }
}

// 0 accept, 1 reject, -1 precondition failure
int orc_plonk_verify(int curve, unsigned lg_n, const u64* C_qs, const u64* pub, size_t n_pub, size_t public_input_count,
                     const u64* acc_prev, const u64* proof, int succinct_only) {
    try { return curve == 0 ? plonk_verify_impl<0>(lg_n, C_qs, pub, n_pub, public_input_count, acc_prev, proof, succinct_only)
                            : plonk_verify_impl<1>(lg_n, C_qs, pub, n_pub, public_input_count, acc_prev, proof, succinct_only); }
    catch (std::exception& e) { g_err = e.what(); return -1; }
}

int orc_acc_decider(int curve, const u64* acc) { return orc_pcdl_check(curve, acc); }

// literal generator fold of one IPA round (pcdl.rs:427-429) on raw affine arrays; for kernel parity tests
void orc_fold_generators(int curve, const u64* gl_xy, const u64* gr_xy, const u64* xi, size_t m, u64* out_xy) {
    if (curve == 0) {
        std::vector<Jac<0>> f(m); Fp x = Fp::from_raw(xi);
        par_for_idx(m, [&](size_t j) { f[j] = Jac<0>::from_affine(Affine<0>::from_raw(gr_xy + 8 * j)).mul(x).add_affine(Affine<0>::from_raw(gl_xy + 8 * j)); });
        auto a = normalize_batch<0>(f); for (size_t j = 0; j < m; ++j) a[j].to_raw(out_xy + 8 * j);
    } else {
        std::vector<Jac<1>> f(m); Fq x = Fq::from_raw(xi);
        par_for_idx(m, [&](size_t j) { f[j] = Jac<1>::from_affine(Affine<1>::from_raw(gr_xy + 8 * j)).mul(x).add_affine(Affine<1>::from_raw(gl_xy + 8 * j)); });
        auto a = normalize_batch<1>(f); for (size_t j = 0; j < m; ++j) a[j].to_raw(out_xy + 8 * j);
    }
}

}  // extern "C"

// ---- ark-serialize compressed encodings (serialize.hpp) ----
extern "C" {
void orc_point_compress(int curve, const u64* xyz, uint8_t* out33) {
    std::vector<uint8_t> b;
    if (curve == 0) ser::write_point<0>(Jac<0>::from_raw(xyz), b); else ser::write_point<1>(Jac<1>::from_raw(xyz), b);
    memcpy(out33, b.data(), 33);
}
int orc_point_decompress(int curve, const uint8_t* in33, u64* out_xyz) {
    if (curve == 0) { Jac<0> p; if (!ser::read_point<0>(in33, &p)) return 1; p.to_raw(out_xyz); }
    else { Jac<1> p; if (!ser::read_point<1>(in33, &p)) return 1; p.to_raw(out_xyz); }
    return 0;
}
// returns the number of bytes written (0 if cap is too small)
size_t orc_instance_serialize(int curve, const u64* inst, uint8_t* out, size_t cap) {
    std::vector<uint8_t> b;
    if (curve == 0) ser::write_instance<0>(instance_from_flat<0>(inst), b); else ser::write_instance<1>(instance_from_flat<1>(inst), b);
    if (b.size() > cap) return 0;
    memcpy(out, b.data(), b.size());
    return b.size();
}
// returns bytes consumed (0 = malformed); *lg_n receives the length of Ls
size_t orc_instance_deserialize(int curve, const uint8_t* in, size_t len, u64* out_inst, size_t cap_words, unsigned* lg_n) {
    size_t used = 0;
    if (curve == 0) {
        Instance<0> q; used = ser::read_instance<0>(in, len, &q);
        if (!used || q.pi.Ls.size() != q.pi.Rs.size() || instance_words((unsigned)q.pi.Ls.size()) > cap_words) return 0;
        if (((size_t)1 << q.pi.Ls.size()) != q.d + 1) return 0;
        instance_to_flat<0>(q, out_inst); *lg_n = (unsigned)q.pi.Ls.size();
    } else {
        Instance<1> q; used = ser::read_instance<1>(in, len, &q);
        if (!used || q.pi.Ls.size() != q.pi.Rs.size() || instance_words((unsigned)q.pi.Ls.size()) > cap_words) return 0;
        if (((size_t)1 << q.pi.Ls.size()) != q.d + 1) return 0;
        instance_to_flat<1>(q, out_inst); *lg_n = (unsigned)q.pi.Ls.size();
    }
    return used;
}
}  // extern "C"
