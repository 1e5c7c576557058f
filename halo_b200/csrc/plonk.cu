// PLONK prover and verifier over the device-resident MSM / NTT / IPA paths.
//
// Stands in for crates/plonk/src/plonk/protocol.rs:64-355 (PlonkProof::naive_prover), :357-502 (verify_succinct,
// verify) and the helpers at :505-1017, for the inputs Trace::consume hands over (crates/plonk/src/circuit/
// trace.rs:137-232): the coefficient vectors of the 10 selector, 16 witness, 15 round-constant, 8 identity and 8
// permutation polynomials, the first three witness columns as evaluations, the public inputs and the previous
// accumulator.  The proof is the same tuple, flattened (include/halo_b200.h).
//
// The reference forms f', g' and the quotient numerator through ~17 FFT-based polynomial products and ~150
// vector passes over the 8n domain.  Here every polynomial stays in HBM and the numerator
//     f = f_GC + alpha * l1 * (z - 1) + alpha^2 * (z f' - z_omega g')            (degree < 9n)
// is evaluated pointwise on the 16 cosets  w16n^c * H  (c = 0..15, H the size-n domain) by one fused kernel per
// coset, fed by one batched size-n coset NTT of all 61 columns; the 16n values are interpolated once and divided by
// X^n - 1 in coefficient form (t_j = sum_{i>=1} f_{j+in}, the exact quotient the reference's
// divide_by_vanishing_poly returns, remainder dropped).  All of these are exact polynomial identities over a
// field, so every commitment, evaluation and opening is bit-identical to the literal formulation.
// The grand product z is two multiplicative scans (prefix of the numerators, suffix of the denominators) and ONE
// field inversion instead of n serial divisions (protocol.rs:144-155).
#include "runtime.cuh"
#include "transcript.cuh"
#include "host.cuh"
#include <algorithm>
#include <cstdio>
#include <vector>

namespace halo {

namespace {

constexpr int NQ = 10, NW = 16, NR = 15, NS = 8, NT = 16;
constexpr int COL_Q = 0, COL_W = 10, COL_R = 26, COL_ID = 41, COL_SIG = 49, COL_PI = 57, COL_Z = 58, COL_ZW = 59,
              COL_L1 = 60, COL_WOM = 61;
constexpr int N_IN = 57;          // columns supplied by the caller
constexpr int N_EVAL = 61;        // columns evaluated on every coset
constexpr int N_COLS = 64;        // all resident coefficient columns

template <int F> HD Fe<F> p_ld(const Fe<F>* p) {
    Fe<F> r;
#if defined(__CUDA_ARCH__)
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
#else
    r = *p;
#endif
    return r;
}
template <int F> __device__ __forceinline__ void p_st(Fe<F>* p, const Fe<F>& v) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    q[1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}

// ---------------------------------------------------------------------------------------------------------
// the gate expressions (protocol.rs:591-1011), written once for the device kernel and the host verifier.
// `V` supplies the values at one point: q(i), w(i), r(i), nw(i) (next row's first three witnesses) and pi().
// Shared factors are pulled out; over a field that changes no value.
// ---------------------------------------------------------------------------------------------------------
template <int F> HD Fe<F> pow7(const Fe<F>& x) {
    Fe<F> x2 = fe_sqr<F>(x), x3 = fe_mul<F>(x2, x);
    return fe_mul<F>(fe_sqr<F>(x3), x);
}

// five Poseidon rounds on one row (protocol.rs:591-625)
template <int F, class V> HD Fe<F> gate_poseidon(const V& v, const Fe<F>* M) {
    Fe<F> acc = fe_zero<F>();
    for (int rd = 0; rd < 5; ++rd) {
        Fe<F> s0 = pow7<F>(v.w(3 * rd)), s1 = pow7<F>(v.w(3 * rd + 1)), s2 = pow7<F>(v.w(3 * rd + 2));
        for (int i = 0; i < 3; ++i) {
            Fe<F> o = rd < 4 ? v.w(3 * rd + 3 + i) : v.nw(i);
            Fe<F> m = fe_add<F>(fe_add<F>(fe_mul<F>(s0, M[3 * i]), fe_mul<F>(s1, M[3 * i + 1])), fe_mul<F>(s2, M[3 * i + 2]));
            acc = fe_add<F>(acc, fe_sub<F>(o, fe_add<F>(v.r(3 * rd + i), m)));
        }
    }
    return acc;
}

// complete addition (x1,y1) + (x2,y2) = (xr,yr) with witnesses al, de, la; e and f are the two identity-case factors
template <int F>
HD Fe<F> add_core(const Fe<F>& x1, const Fe<F>& y1, const Fe<F>& x2, const Fe<F>& y2, const Fe<F>& xr, const Fe<F>& yr,
                  const Fe<F>& al, const Fe<F>& de, const Fe<F>& la, const Fe<F>& e, const Fe<F>& f) {
    const Fe<F> one = fe_one<F>();
    Fe<F> dx = fe_sub<F>(x2, x1), dy = fe_sub<F>(y2, y1), sy = fe_add<F>(y2, y1);
    Fe<F> res = fe_mul<F>(dx, fe_sub<F>(fe_mul<F>(dx, la), dy));
    Fe<F> x1s = fe_sqr<F>(x1);
    Fe<F> dxal = fe_mul<F>(dx, al);
    res = fe_add<F>(res, fe_mul<F>(fe_sub<F>(one, dxal),
                                   fe_sub<F>(fe_mul<F>(fe_dbl<F>(y1), la), fe_add<F>(fe_dbl<F>(x1s), x1s))));
    Fe<F> x1x2 = fe_mul<F>(x1, x2);
    Fe<F> AD = fe_mul<F>(x1x2, fe_add<F>(dx, sy));
    Fe<F> Bv = fe_sub<F>(fe_sub<F>(fe_sub<F>(fe_sqr<F>(la), x1), x2), xr);
    Fe<F> Cv = fe_sub<F>(fe_sub<F>(fe_mul<F>(la, fe_sub<F>(x1, xr)), y1), yr);
    res = fe_add<F>(res, fe_mul<F>(AD, fe_add<F>(Bv, Cv)));
    res = fe_add<F>(res, fe_mul<F>(e, fe_add<F>(fe_sub<F>(xr, x2), fe_sub<F>(yr, y2))));
    res = fe_add<F>(res, fe_mul<F>(f, fe_add<F>(fe_sub<F>(xr, x1), fe_sub<F>(yr, y1))));
    Fe<F> g = fe_sub<F>(fe_sub<F>(one, dxal), fe_mul<F>(sy, de));
    return fe_add<F>(res, fe_mul<F>(g, fe_add<F>(xr, yr)));
}

// protocol.rs:650-703: columns xp yp xq yq xr yr alpha beta gamma delta lambda
template <int F, class V> HD Fe<F> gate_affine_add(const V& v) {
    const Fe<F> one = fe_one<F>();
    Fe<F> xp = v.w(0), xq = v.w(2);
    Fe<F> e = fe_sub<F>(one, fe_mul<F>(xp, v.w(7))), f = fe_sub<F>(one, fe_mul<F>(xq, v.w(8)));
    return add_core<F>(xp, v.w(1), xq, v.w(3), v.w(4), v.w(5), v.w(6), v.w(9), v.w(10), e, f);
}

// protocol.rs:763-849: one double-and-add step, columns xp yp a xg yg b xq yq xr yr bq lq ar gr dr lr
template <int F, class V> HD Fe<F> gate_affine_mul(const V& v) {
    const Fe<F> one = fe_one<F>();
    Fe<F> xp = v.w(0), yp = v.w(1), a = v.w(2), xg = v.w(3), yg = v.w(4), b = v.w(5), xq = v.w(6), yq = v.w(7),
          xr = v.w(8), yr = v.w(9), lq = v.w(11);
    Fe<F> e = fe_sub<F>(one, fe_mul<F>(xp, v.w(10)));
    Fe<F> xps = fe_sqr<F>(xp);
    Fe<F> res = fe_mul<F>(e, fe_add<F>(xq, yq));
    res = fe_add<F>(res, fe_sub<F>(fe_mul<F>(fe_dbl<F>(yp), lq), fe_add<F>(fe_dbl<F>(xps), xps)));
    res = fe_add<F>(res, fe_sub<F>(fe_sub<F>(fe_sqr<F>(lq), fe_dbl<F>(xp)), xq));
    res = fe_add<F>(res, fe_sub<F>(fe_sub<F>(fe_mul<F>(lq, fe_sub<F>(xp, xq)), yp), yq));
    Fe<F> f = fe_sub<F>(one, fe_mul<F>(xg, v.w(13)));
    res = fe_add<F>(res, add_core<F>(xq, yq, xg, yg, xr, yr, v.w(12), v.w(14), v.w(15), e, f));
    res = fe_add<F>(res, fe_mul<F>(b, fe_sub<F>(b, one)));
    Fe<F> nb = fe_sub<F>(one, b);
    res = fe_add<F>(res, fe_sub<F>(v.nw(0), fe_add<F>(fe_mul<F>(b, xr), fe_mul<F>(nb, xq))));
    res = fe_add<F>(res, fe_sub<F>(v.nw(1), fe_add<F>(fe_mul<F>(b, yr), fe_mul<F>(nb, yq))));
    return fe_sub<F>(fe_add<F>(res, v.nw(2)), fe_add<F>(a, fe_mul<F>(b, v.r(0))));
}

// protocol.rs:939-964
template <int F, class V> HD Fe<F> gate_eq(const V& v) {
    Fe<F> d = fe_sub<F>(v.w(0), v.w(1)), eq = v.w(3);
    return fe_add<F>(fe_mul<F>(d, fe_add<F>(eq, v.w(4))), fe_sub<F>(eq, v.w(2)));
}

// protocol.rs:851-937
template <int F, class V> HD Fe<F> gate_range(const V& v) {
    Fe<F> res = fe_sub<F>(v.nw(0), v.w(0));
    for (int i = 0; i < 15; ++i) res = fe_sub<F>(res, fe_mul<F>(v.w(1 + i), v.r(i)));
    return res;
}

// f_GC at one point (protocol.rs:183-193); gate_mask bit i = selector q_{5+i} is not the zero polynomial
template <int F, class V> HD Fe<F> f_gc_point(const V& v, const Fe<F>* M, unsigned gate_mask) {
    Fe<F> w0 = v.w(0), w1 = v.w(1);
    Fe<F> res = fe_add<F>(fe_mul<F>(w0, v.q(0)), fe_mul<F>(v.q(1), w1));
    res = fe_add<F>(res, fe_mul<F>(v.q(2), v.w(2)));
    res = fe_add<F>(res, fe_mul<F>(fe_mul<F>(v.q(3), w0), w1));
    res = fe_add<F>(res, v.q(4));
    if (gate_mask & 1u) res = fe_add<F>(res, fe_mul<F>(v.q(5), gate_poseidon<F, V>(v, M)));
    if (gate_mask & 2u) res = fe_add<F>(res, fe_mul<F>(v.q(6), gate_affine_add<F, V>(v)));
    if (gate_mask & 4u) res = fe_add<F>(res, fe_mul<F>(v.q(7), gate_affine_mul<F, V>(v)));
    if (gate_mask & 8u) res = fe_add<F>(res, fe_mul<F>(v.q(8), gate_eq<F, V>(v)));
    if (gate_mask & 16u) res = fe_add<F>(res, fe_mul<F>(v.q(9), gate_range<F, V>(v)));
    return fe_add<F>(res, v.pi());
}

// ---------------------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------------------
template <int F> struct Challenges { Fe<F> beta, gamma, alpha, alpha2, M[9], id_step_beta; };

// Columns that are the zero polynomial (unused selectors, witness and round-constant columns of a circuit) are not
// transformed at all: slot[c] is the position of column c among the transformed ones, and every zero column reads
// one shared column of zeros.
struct ColMap { unsigned char slot[N_COLS]; };

// values of the 61 columns at point j of the current coset; the omega-shift is index j+1 in the same coset
template <int F> struct ColView {
    const Fe<F>* ev; size_t n, j, jn; const ColMap& m;
    HD Fe<F> col(int c) const { return p_ld<F>(ev + (size_t)m.slot[c] * n + j); }
    HD Fe<F> q(int i) const { return col(COL_Q + i); }
    HD Fe<F> w(int i) const { return col(COL_W + i); }
    HD Fe<F> r(int i) const { return col(COL_R + i); }
    HD Fe<F> nw(int i) const { return p_ld<F>(ev + (size_t)m.slot[COL_W + i] * n + jn); }
    HD Fe<F> pi() const { return col(COL_PI); }
};

// f on one coset: out[coset + stride * j] = f(g * w_n^j) [* scale]
template <int F>
__global__ void __launch_bounds__(128) plonk_numerator_kernel(const Fe<F>* __restrict__ ev, size_t n, Challenges<F> ch,
                                                              unsigned gate_mask, Fe<F>* __restrict__ out,
                                                              unsigned coset, unsigned stride, Fe<F> scale, int scaled,
                                                              const __grid_constant__ ColMap map) {
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n) return;
    ColView<F> v{ev, n, j, j + 1 == n ? 0 : j + 1, map};
    Fe<F> f = f_gc_point<F, ColView<F>>(v, ch.M, gate_mask);
    // f' and g' (protocol.rs:132-139)
    // id_step != 0: the identity columns differ by a constant (id_i = id_0 + i * id_step, the reference's
    // SlotId::to_scalar numbering), so only id_0 was transformed and beta * id_i is a running sum
    Fe<F> fp = fe_one<F>(), gp = fe_one<F>();
    const bool ids_derived = !fe_is_zero<F>(ch.id_step_beta);
    Fe<F> idb = fe_mul<F>(v.col(COL_ID), ch.beta);
    for (int i = 0; i < NS; ++i) {
        Fe<F> wg = fe_add<F>(v.w(i), ch.gamma);
        if (i) idb = ids_derived ? fe_add<F>(idb, ch.id_step_beta) : fe_mul<F>(v.col(COL_ID + i), ch.beta);
        fp = fe_mul<F>(fp, fe_add<F>(wg, idb));
        gp = fe_mul<F>(gp, fe_add<F>(wg, fe_mul<F>(v.col(COL_SIG + i), ch.beta)));
    }
    Fe<F> z = v.col(COL_Z);
    Fe<F> zw = p_ld<F>(ev + (size_t)map.slot[COL_Z] * n + v.jn);      // z_omega(x) = z(omega x): the next point of the coset
    Fe<F> cc1 = fe_mul<F>(v.col(COL_L1), fe_sub<F>(z, fe_one<F>()));
    Fe<F> cc2 = fe_sub<F>(fe_mul<F>(z, fp), fe_mul<F>(zw, gp));
    f = fe_add<F>(f, fe_add<F>(fe_mul<F>(ch.alpha, cc1), fe_mul<F>(ch.alpha2, cc2)));
    if (scaled) f = fe_mul<F>(f, scale);                  // 1 / Z_H on this coset (quotient path)
    p_st<F>(out + coset + (size_t)stride * j, f);
}

// numerators and denominators of the grand product on H (protocol.rs:140-141): ev holds w_0..7, id_0..7, sigma_0..7
// map.slot[i], [NS + i], [2 NS + i]: where w_i, id_i, sigma_i sit among the transformed columns (zero columns share a
// column of zeros); id_step_beta != 0: only id_0 was transformed, beta * id_i = beta * id_0 + i * id_step_beta
template <int F>
__global__ void __launch_bounds__(128) plonk_perm_kernel(const Fe<F>* __restrict__ ev, size_t n, Fe<F> beta, Fe<F> gamma,
                                                         Fe<F>* __restrict__ num, Fe<F>* __restrict__ den,
                                                         Fe<F> id_step_beta, const __grid_constant__ ColMap map) {
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n) return;
    Fe<F> fp = fe_one<F>(), gp = fe_one<F>();
    if (j) {                                              // index 0 is not part of the running product
        const bool ids_derived = !fe_is_zero<F>(id_step_beta);
        Fe<F> idb = fe_mul<F>(p_ld<F>(ev + (size_t)map.slot[NS] * n + j), beta);
        for (int i = 0; i < NS; ++i) {
            Fe<F> wg = fe_add<F>(p_ld<F>(ev + (size_t)map.slot[i] * n + j), gamma);
            if (i) idb = ids_derived ? fe_add<F>(idb, id_step_beta) : fe_mul<F>(p_ld<F>(ev + (size_t)map.slot[NS + i] * n + j), beta);
            fp = fe_mul<F>(fp, fe_add<F>(wg, idb));
            gp = fe_mul<F>(gp, fe_add<F>(wg, fe_mul<F>(p_ld<F>(ev + (size_t)map.slot[2 * NS + i] * n + j), beta)));
        }
    }
    p_st<F>(num + j, fp);
    p_st<F>(den + j, gp);
}

// multiplicative scan, in place: a[i] <- a[0] * ... * a[i]  (reverse: a[i] <- a[i] * ... * a[n-1]).
// One CTA scans SCAN_SEG consecutive elements and leaves its total in totals[block].
constexpr int SCAN_T = 256, SCAN_E = 8, SCAN_SEG = SCAN_T * SCAN_E;
template <int F>
__global__ void __launch_bounds__(SCAN_T) scan_local_kernel(Fe<F>* __restrict__ a, size_t n, Fe<F>* __restrict__ totals,
                                                            int reverse) {
    __shared__ uint4 sh[2 * SCAN_T];
    const size_t base = (size_t)blockIdx.x * SCAN_SEG + (size_t)threadIdx.x * SCAN_E;
    Fe<F> p[SCAN_E];
    Fe<F> run = fe_one<F>();
#pragma unroll
    for (int e = 0; e < SCAN_E; ++e) {
        size_t L = base + e;
        if (L < n) run = fe_mul<F>(run, p_ld<F>(a + (reverse ? n - 1 - L : L)));
        p[e] = run;
    }
    auto put = [&](int t, const Fe<F>& v) {
        sh[2 * t] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
        sh[2 * t + 1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
    };
    auto get = [&](int t) {
        Fe<F> r; uint4 x = sh[2 * t], y = sh[2 * t + 1];
        r.l[0] = x.x; r.l[1] = x.y; r.l[2] = x.z; r.l[3] = x.w; r.l[4] = y.x; r.l[5] = y.y; r.l[6] = y.z; r.l[7] = y.w;
        return r;
    };
    const int t = threadIdx.x;
    put(t, run);
    __syncthreads();
    for (int off = 1; off < SCAN_T; off <<= 1) {          // Hillis-Steele over the thread totals
        Fe<F> other = fe_one<F>();
        if (t >= off) other = get(t - off);
        __syncthreads();
        if (t >= off) { run = fe_mul<F>(run, other); put(t, run); }
        __syncthreads();
    }
    Fe<F> pre = t ? get(t - 1) : fe_one<F>();
#pragma unroll
    for (int e = 0; e < SCAN_E; ++e) {
        size_t L = base + e;
        if (L < n) p_st<F>(a + (reverse ? n - 1 - L : L), t ? fe_mul<F>(p[e], pre) : p[e]);
    }
    if (t == SCAN_T - 1) p_st<F>(totals + blockIdx.x, run);
}
// second phase: multiply segment b > 0 by the product of all earlier segments (scanned totals)
template <int F>
__global__ void __launch_bounds__(256) scan_apply_kernel(Fe<F>* __restrict__ a, size_t n, const Fe<F>* __restrict__ totals,
                                                         int reverse) {
    size_t L = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (L >= n || L < SCAN_SEG) return;
    Fe<F> pre = p_ld<F>(totals + L / SCAN_SEG - 1);
    Fe<F>* p = a + (reverse ? n - 1 - L : L);
    p_st<F>(p, fe_mul<F>(p_ld<F>(p), pre));
}

// z in the Evals convention and its omega-shift (protocol.rs:144-159): zv[i] = P[i] * S[i+1] / T with P the prefix
// products of the numerators, S the suffix products of the denominators and T = S[0]; z_evals[(i+1) % n] = zv[i]
template <int F>
__global__ void __launch_bounds__(256) plonk_z_kernel(const Fe<F>* __restrict__ P, const Fe<F>* __restrict__ S, size_t n,
                                                      Fe<F> inv_total, Fe<F>* __restrict__ z_evals, Fe<F>* __restrict__ zw_evals) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fe<F> v = fe_mul<F>(p_ld<F>(P + i), inv_total);
    if (i + 1 < n) v = fe_mul<F>(v, p_ld<F>(S + i + 1));
    p_st<F>(z_evals + (i + 1 == n ? 0 : i + 1), v);
    p_st<F>(zw_evals + i, v);
}

// quotient by X^n - 1 in coefficient form: t[j] = sum_{i>=1, j+in < len} f[j + i n]   (ark-poly's
// divide_by_vanishing_poly; the remainder is dropped as at protocol.rs:256)
template <int F>
__global__ void __launch_bounds__(256) plonk_quotient_kernel(const Fe<F>* __restrict__ f, size_t len, size_t n,
                                                             Fe<F>* __restrict__ t) {
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= len) return;
    Fe<F> acc = fe_zero<F>();
    for (size_t k = j + n; k < len; k += n) acc = fe_add<F>(acc, p_ld<F>(f + k));
    p_st<F>(t + j, acc);
}

// out[k] = sum_i coef[i] * col_i[k]   (geometric_polys, protocol.rs:542-548)
constexpr int LC_MAX = 44;
template <int F> struct LinComb { const Fe<F>* col[LC_MAX]; Fe<F> coef[LC_MAX]; int count; };
template <int F>
__global__ void __launch_bounds__(256) plonk_lincomb_kernel(LinComb<F> lc, size_t n, Fe<F>* __restrict__ out) {
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    Fe<F> acc = p_ld<F>(lc.col[0] + k);                   // coef[0] = 1
    for (int i = 1; i < lc.count; ++i) acc = fe_add<F>(acc, fe_mul<F>(lc.coef[i], p_ld<F>(lc.col[i] + k)));
    p_st<F>(out + k, acc);
}

// pw[k] = x^k, 16 consecutive powers per thread
template <int F>
__global__ void __launch_bounds__(128) plonk_powers_kernel(Fe<F>* __restrict__ pw, size_t n, Fe<F> x) {
    size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t lo = c * 16;
    if (lo >= n) return;
    Fe<F> cur = fe_one<F>(), b = x;
    for (size_t e = lo; e; e >>= 1) { if (e & 1) cur = fe_mul<F>(cur, b); b = fe_sqr<F>(b); }
    for (size_t k = lo; k < lo + 16 && k < n; ++k) { p_st<F>(pw + k, cur); cur = fe_mul<F>(cur, x); }
}

// partial[y * gridDim.x + x] = sum over this block's slice of col_y[k] * pw_y[k]
constexpr int EV_MAX = 80, EV_BLOCKS = 64;
template <int F> struct EvalSet { const Fe<F>* col[EV_MAX]; const Fe<F>* pw[EV_MAX]; };
template <int F>
__global__ void __launch_bounds__(256) plonk_eval_kernel(EvalSet<F> es, size_t n, Fe<F>* __restrict__ partial) {
    __shared__ uint4 sh[2 * 256];
    const Fe<F>* c = es.col[blockIdx.y];
    const Fe<F>* pw = es.pw[blockIdx.y];
    Fe<F> acc = fe_zero<F>();
    for (size_t k = (size_t)blockIdx.x * 256 + threadIdx.x; k < n; k += (size_t)gridDim.x * 256)
        acc = fe_add<F>(acc, fe_mul<F>(p_ld<F>(c + k), p_ld<F>(pw + k)));
    const int t = threadIdx.x;
    sh[2 * t] = make_uint4(acc.l[0], acc.l[1], acc.l[2], acc.l[3]);
    sh[2 * t + 1] = make_uint4(acc.l[4], acc.l[5], acc.l[6], acc.l[7]);
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (t < off) {
            Fe<F> o; uint4 x = sh[2 * (t + off)], y = sh[2 * (t + off) + 1];
            o.l[0] = x.x; o.l[1] = x.y; o.l[2] = x.z; o.l[3] = x.w; o.l[4] = y.x; o.l[5] = y.y; o.l[6] = y.z; o.l[7] = y.w;
            acc = fe_add<F>(acc, o);
            sh[2 * t] = make_uint4(acc.l[0], acc.l[1], acc.l[2], acc.l[3]);
            sh[2 * t + 1] = make_uint4(acc.l[4], acc.l[5], acc.l[6], acc.l[7]);
        }
        __syncthreads();
    }
    if (t == 0) p_st<F>(partial + (size_t)blockIdx.y * gridDim.x + blockIdx.x, acc);
}

// dst[c][k] = src[list[c]][k] * g^k for the ncols listed columns: the coset shift of a batched size-n NTT, fused with
// the (compacting) copy into the transform's scratch buffer.  g^k = hi[k >> 10] * lo[k & 1023]; one thread per k,
// coalesced across columns.
template <int F>
__global__ void __launch_bounds__(256) plonk_coset_stage_kernel(const Fe<F>* __restrict__ src, Fe<F>* __restrict__ dst, size_t n,
                                                                int ncols, const Fe<F>* __restrict__ lo,
                                                                const Fe<F>* __restrict__ hi,
                                                                const __grid_constant__ ColMap list) {
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    Fe<F> w = p_ld<F>(lo + (k & 1023));
    if (k >> 10) w = fe_mul<F>(w, p_ld<F>(hi + (k >> 10)));
    for (int c = 0; c < ncols; ++c)
        p_st<F>(dst + (size_t)c * n + k, fe_mul<F>(p_ld<F>(src + (size_t)list.slot[c] * n + k), w));
}

// any nonzero coefficient in column c?  flags[c] |= 1
template <int F>
__global__ void __launch_bounds__(256) plonk_nonzero_kernel(const Fe<F>* __restrict__ cols, size_t n, int* __restrict__ flags) {
    const Fe<F>* c = cols + (size_t)blockIdx.y * n;
    u32 any = 0;
    for (size_t k = (size_t)blockIdx.x * 256 + threadIdx.x; k < n; k += (size_t)gridDim.x * 256) {
        Fe<F> v = p_ld<F>(c + k);
        for (int i = 0; i < 8; ++i) any |= v.l[i];
    }
    if (__syncthreads_or(any != 0) && threadIdx.x == 0) atomicOr(flags + blockIdx.y, 1);
}

// do the identity columns differ by a constant?  flags[0] |= 1 unless id_i - id_0 == i * step (a constant polynomial)
// for every i = 1..7 (blockIdx.y + 1)
template <int F>
__global__ void __launch_bounds__(256) plonk_id_affine_kernel(const Fe<F>* __restrict__ ids, size_t n, Fe<F> step,
                                                              int* __restrict__ flags) {
    const int i = blockIdx.y + 1;
    Fe<F> want = fe_zero<F>();
    for (int t = 0; t < i; ++t) want = fe_add<F>(want, step);
    u32 bad = 0;
    for (size_t k = (size_t)blockIdx.x * 256 + threadIdx.x; k < n; k += (size_t)gridDim.x * 256) {
        Fe<F> d = fe_sub<F>(p_ld<F>(ids + (size_t)i * n + k), p_ld<F>(ids + k));
        bad |= k == 0 ? !fe_eq<F>(d, want) : !fe_is_zero<F>(d);
    }
    if (__syncthreads_or(bad != 0) && threadIdx.x == 0) atomicOr(flags, 1);
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
template <int F> Fe<F> root_of_unity(int log_n) {
    // 5^((p-1)/2^32), squared down to order 2^log_n  (ark-ff: GENERATOR = 5, TWO_ADICITY = 32)
    u32 e[8];
    for (int i = 0; i < 7; ++i) e[i] = modulus_limb<F>(i + 1);
    e[7] = 0;
    Fe<F> w = fe_pow<F>(fe_from_u64<F>(5), e, 256);
    for (int i = log_n; i < 32; ++i) w = fe_sqr<F>(w);
    return w;
}

template <int F> int scan_mul_device(Fe<F>* a, size_t n, Fe<F>* scratch, bool reverse) {
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    const size_t nblk = (n + SCAN_SEG - 1) / SCAN_SEG;
    HALO_CHECK(nblk <= (size_t)SCAN_SEG, HALO_EINVAL, "plonk: scan length above 2^22");
    scan_local_kernel<F><<<(unsigned)nblk, SCAN_T, 0, st>>>(a, n, scratch, reverse ? 1 : 0);
    R.launches += 1;
    if (nblk > 1) {
        scan_local_kernel<F><<<1, SCAN_T, 0, st>>>(scratch, nblk, scratch + nblk, 0);
        scan_apply_kernel<F><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a, n, scratch, reverse ? 1 : 0);
        R.launches += 2;
    }
    HALO_CUDA(cudaGetLastError());
    return HALO_OK;
}

static inline size_t evalproof_words(unsigned lg_n) { return 33 + 24 * (size_t)lg_n; }
static inline size_t instance_words(unsigned lg_n) { return 54 + 24 * (size_t)lg_n; }

}  // namespace

// out[j * K + k] = in[k * n + j]: K coset-major blocks of n values -> the interleaved order of the K n-point transform
template <int F>
__global__ void __launch_bounds__(256) plonk_interleave_kernel(const Fe<F>* __restrict__ in, size_t n, unsigned K, Fe<F>* __restrict__ out) {
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= n * K) return;
    const size_t j = t / K; const unsigned k = (unsigned)(t - j * K);
    p_st<F>(out + t, p_ld<F>(in + (size_t)k * n + j));
}

size_t plonk_proof_words(unsigned lg_n) { return 78 * 4 + 33 * 12 + 2 * evalproof_words(lg_n) + instance_words(lg_n); }

// proof layout (u64 words): vs = ws[16] rs[15] qs[10] ts[16] ids[8] sigmas[8] z z_omega w_omegas[3] (4 each),
// Cs = ws[16] ts[16] z (12 each), pi_r, pi_r_omega (EvalProof: Ls Rs U c hiding C_bar w_prime), acc_next (instance)
// d_circuit == nullptr: polys holds all 57 columns.  Otherwise polys holds the 16 witness columns only and d_circuit
// the 41 circuit columns (q r id sigma) already on the device (plonk_circuit_load).
// shard == true (halo_dist_plonk_prove): ONE proof over the GPUs of the library's communicator.  Every rank runs the
// same control flow and the same transcript; the heavy steps are divided: rank 0 uploads and broadcasts the columns, the
// commitments are point-range-sharded MSMs, the cosets of the quotient are divided among the ranks and all-gathered, the
// two openings run on ranks 0 and 1; the serial pieces (grand product, evaluations, accumulation) are replicated.
// polys / w_evals are read on rank 0 only.
template <int C>
int plonk_prove(unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* pub, size_t n_pub,
                const uint64_t* acc_prev, uint64_t* out, const void* d_circuit, bool shard, uint64_t* residual_out) {
    // residual_out != nullptr (halo_plonk_gate_residual): stop at round 4 and return f_GC on H instead of a proof
    typedef Host<C> H;
    typedef typename H::SF SF;
    constexpr int S = H::S;
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    const bool sharded = shard && dist_on() && dist_world() > 1;
    const int rank = sharded ? dist_rank() : 0, world = sharded ? dist_world() : 1;
    HALO_CHECK(lg_n >= 1 && lg_n <= 22, HALO_EINVAL, "plonk: rows must be a power of two in [2, 2^22]");
    const size_t n = (size_t)1 << lg_n, d = n - 1;
    HALO_CHECK(R.pp[C].n >= n, HALO_EPARAMS, "plonk: public parameters not loaded for this many rows");
    HALO_CHECK(n_pub <= n, HALO_EINVAL, "plonk: more public inputs than rows");
    HALO_CHECK(acc_prev[12] == d, HALO_EINVAL, "plonk: acc_prev was made for another size");
    const bool trace = R.tune[7] == 1;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (trace) { cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventRecord(ev0, st); }
    auto lap = [&](const char* what) {
        if (!trace) return;
        cudaEventRecord(ev1, st); cudaEventSynchronize(ev1);
        float ms = 0; cudaEventElapsedTime(&ms, ev0, ev1);
        fprintf(stderr, "[halo trace] plonk %-28s %9.3f ms\n", what, ms);
        cudaEventRecord(ev0, st);
    };

    SF *cols, *tmp, *ev, *f0, *f1, *misc;
    if (int e = ws_get(WS_PLONK_COLS, (size_t)N_COLS * n * 32, (void**)&cols)) return e;
    if (int e = ws_get(WS_PLONK_TMP, (size_t)N_EVAL * n * 32, (void**)&tmp)) return e;
    if (int e = ws_get(WS_PLONK_EV, (size_t)N_EVAL * n * 32, (void**)&ev)) return e;
    if (int e = ws_get(WS_PLONK_F0, 16 * n * 32, (void**)&f0)) return e;
    if (int e = ws_get(WS_PLONK_F1, 16 * n * 32, (void**)&f1)) return e;
    const size_t gtab_elems = 1024 + n / 1024 + 16;
    const size_t misc_elems = 2 * (size_t)SCAN_SEG + 64 + (size_t)EV_MAX * EV_BLOCKS + gtab_elems;
    if (int e = ws_get(WS_PLONK_MISC, misc_elems * 32 + 256, (void**)&misc)) return e;
    int* d_flags = (int*)(misc + misc_elems);
    SF* gtab = misc + 2 * (size_t)SCAN_SEG + 64 + (size_t)EV_MAX * EV_BLOCKS;      // g^i, i < 1024 | (g^1024)^i
    auto col = [&](int c) { return cols + (size_t)c * n; };
    const SF one = fe_one<S>();
    const SF omega = root_of_unity<S>((int)lg_n);

    // ---- round 0 (protocol.rs:74-109): resident coefficient columns
    if (d_circuit) {
        const SF* dc = (const SF*)d_circuit;
        HALO_CUDA(cudaMemcpyAsync(col(COL_Q), dc, (size_t)NQ * n * 32, cudaMemcpyDeviceToDevice, st));
        HALO_CUDA(cudaMemcpyAsync(col(COL_R), dc + (size_t)NQ * n, (size_t)(NR + 2 * NS) * n * 32, cudaMemcpyDeviceToDevice, st));
    }
    // staged evaluation vectors: PI (row i at omega^(i+1), negated), l1, and the three shifted witness columns
    SF* stage = tmp;                                       // [PI | l1 | wom0 wom1 wom2]
    // Round 1 needs the witness columns only, so they are uploaded FIRST and (one GPU) their commitments start at once
    // on a helper thread in the library's second execution context, beside the upload of the other 41 columns, the staged
    // evaluation vectors and the small inverse transforms -- the GPU is otherwise idle while the host stages 1.4 GB.
    // Tuning key 15 = 1: commitments after round 0 as before.
    const bool dense = R.tune[6] == 1;
    int h_flags[N_COLS];
    for (int i = 0; i < N_COLS; ++i) h_flags[i] = 0;
    auto nonzero = [&](int c) { return dense || h_flags[c] != 0; };
    // nb adjacent columns -> nb commitments as one batched MSM over the resident generators; sharded: every rank takes
    // its point range of every column (packed into a compact batch) and the window sums are all-gathered (dist.cu)
    const size_t pr_lo = n * (size_t)rank / (size_t)world, pr_hi = n * (size_t)(rank + 1) / (size_t)world;
    auto commit_run = [&](const SF* first, size_t nb, uint64_t* o) -> int {
        Runtime& Rk = rt();                                 // the main runtime, or the lane inside a helper thread
        if (!sharded) return msm_device<C>(Rk.pp[C].d_gs, first, n, nb, o, 0);
        const size_t kl = pr_hi - pr_lo;
        SF* pack = nullptr;
        if (kl) {
            if (int e = ws_get(WS_PLONK_PACK, nb * kl * 32, (void**)&pack)) return e;
            HALO_CUDA(cudaMemcpy2DAsync(pack, kl * 32, first + pr_lo, n * 32, kl * 32, nb, cudaMemcpyDeviceToDevice, st));
        }
        return dist_msm_device<C>((const char*)Rk.pp[C].d_gs + pr_lo * 64, pack, kl, nb, o);
    };
    // commitments of `count` adjacent columns: one batched MSM per run of nonzero columns
    auto commit_cols = [&](const SF* first, int count, auto&& live, uint64_t* o) -> int {
        const Jac<C> id = xyzz_to_jac<C>(xyzz_identity<C>());
        for (int i = 0; i < count;) {
            if (!live(i)) { memcpy(o + 12 * i, &id, 96); ++i; continue; }
            int j = i + 1;
            while (j < count && live(j)) ++j;
            if (int e = commit_run(first + (size_t)i * n, (size_t)(j - i), o + 12 * i)) return e;
            i = j;
        }
        return HALO_OK;
    };
    uint64_t* o_vs = out;
    uint64_t* o_Cws = out + 78 * 4;
    uint64_t* o_Cts = o_Cws + NW * 12;
    uint64_t* o_Cz = o_Cts + NT * 12;
    uint64_t* o_pir = o_Cz + 12;
    uint64_t* o_piro = o_pir + evalproof_words(lg_n);
    uint64_t* o_acc = o_piro + evalproof_words(lg_n);
    Runtime* lane1 = nullptr;
    int rc_w = HALO_OK;
    std::string err_w;
    Workers helper1;
    if (rank == 0) {
        const uint64_t* w_src = d_circuit ? polys : polys + (size_t)COL_W * n * 4;
        if (int e = upload_staged(col(COL_W), w_src, (size_t)NW * n * 32)) return e;
        if (!sharded && !R.is_lane && R.tune[15] != 1 && n >= 4096) {
            HALO_CUDA(cudaMemsetAsync(d_flags + COL_W, 0, NW * sizeof(int), st));
            plonk_nonzero_kernel<S><<<dim3(32, NW), 256, 0, st>>>(col(COL_W), n, d_flags + COL_W);
            R.launches += 1;
            HALO_CUDA(cudaMemcpyAsync(h_flags + COL_W, d_flags + COL_W, NW * sizeof(int), cudaMemcpyDeviceToHost, st));
            HALO_CUDA(cudaStreamSynchronize(st));          // the witness columns are in place, their flags on the host
            if (R.tune[14] != 1 && 2 * n >= R.pp[C].n) {    // the commit-path table is built by the main runtime only
                const void* t_ = nullptr; int c_ = 0, k_ = 0; size_t s_ = 0;
                if (int e = ensure_fixed_table<C>(&t_, &c_, &k_, &s_)) return e;
            }
            lane1 = lane_get();
            if (lane1) {
                helper1.spawn(1, [&]() {
                    LaneScope scope(lane1);
                    try {
                        rc_w = cudaSetDevice(lane1->device) == cudaSuccess
                                   ? commit_cols(col(COL_W), NW, [&](int i) { return nonzero(COL_W + i); }, o_Cws) : HALO_ECUDA;
                    } catch (...) { rc_w = HALO_ECUDA; last_err() = "plonk: exception in the witness commitments"; }
                    if (rc_w) err_w = last_err();
                    cudaStreamSynchronize(lane1->stream);
                });
                if (helper1.size() == 0) lane1 = nullptr;
            }
        }
        if (!d_circuit) {                                  // the other columns: [0, COL_W) and [COL_W + NW, N_IN)
            if (int e = upload_staged(cols, polys, (size_t)COL_W * n * 32)) return e;
            if (int e = upload_staged(col(COL_W + NW), polys + (size_t)(COL_W + NW) * n * 4, (size_t)(N_IN - COL_W - NW) * n * 32)) return e;
        }
        HALO_CUDA(cudaMemsetAsync(stage, 0, 2 * n * 32, st));
        std::vector<SF> neg_pub(n_pub);
        for (size_t i = 0; i < n_pub; ++i) neg_pub[i] = fe_neg<S>(H::scalar(pub + 4 * i));
        if (n_pub) {
            size_t head = std::min(n_pub, n - 1);
            HALO_CUDA(cudaMemcpyAsync(stage + 1, neg_pub.data(), head * 32, cudaMemcpyHostToDevice, st));
            if (n_pub == n) HALO_CUDA(cudaMemcpyAsync(stage, neg_pub.data() + (n - 1), 32, cudaMemcpyHostToDevice, st));
        }
        HALO_CUDA(cudaMemcpyAsync(stage + n + (n > 1 ? 1 : 0), &one, 32, cudaMemcpyHostToDevice, st));
        for (int i = 0; i < 3; ++i) {                      // shift_left of w_evals[i]
            const uint64_t* src = w_evals + (size_t)i * n * 4;
            SF* dst = stage + (size_t)(2 + i) * n;
            if (int e = upload_staged(dst, src + 4, (n - 1) * 32)) return e;
            HALO_CUDA(cudaMemcpyAsync(dst + (n - 1), src, 32, cudaMemcpyHostToDevice, st));
        }
        HALO_CUDA(cudaStreamSynchronize(st));              // neg_pub / one are host temporaries
    }
    if (sharded) {                                         // the columns cross the host link once and NVLink world - 1 times
        if (d_circuit) { if (int e = dist_bcast_dev(col(COL_W), (size_t)NW * n * 32, 0)) return e; }
        else if (int e = dist_bcast_dev(cols, (size_t)N_IN * n * 32, 0)) return e;
        if (int e = dist_bcast_dev(stage, (size_t)5 * n * 32, 0)) return e;
    }
    if (int e = ntt_device<S>(stage, col(COL_PI), (int)lg_n, 1, true, nullptr)) return e;
    if (int e = ntt_device<S>(stage + n, col(COL_L1), (int)lg_n, 4, true, nullptr)) return e;   // l1, wom0..2 are adjacent
    // which columns are the zero polynomial: their commitments are the identity, their coset evaluations are zero,
    // and the custom gates of an all-zero selector are switched off (tuning key 6 = 1 treats every column as dense)
    HALO_CUDA(cudaMemsetAsync(d_flags, 0, N_COLS * sizeof(int), st));
    plonk_nonzero_kernel<S><<<dim3(32, N_IN), 256, 0, st>>>(cols, n, d_flags);
    R.launches += 1;
    // the reference numbers the slots row + (column - 1) * n (SlotId::to_scalar), i.e. id_i = id_0 + i * n as polynomials;
    // verified here on the caller's columns rather than assumed (d_flags[N_COLS - 1] is free until round 3)
    const SF id_step = fe_from_u64<S>((u64)n);
    plonk_id_affine_kernel<S><<<dim3(32, NS - 1), 256, 0, st>>>(col(COL_ID), n, id_step, d_flags + (N_COLS - 1));
    R.launches += 1;
    int ids_not_affine = 1;
    {
        // (a helper thread may be reading the witness entries of h_flags: they are rewritten with the same values only
        // when no helper runs)
        int f_all[N_IN];
        HALO_CUDA(cudaMemcpyAsync(f_all, d_flags, N_IN * sizeof(int), cudaMemcpyDeviceToHost, st));
        HALO_CUDA(cudaMemcpyAsync(&ids_not_affine, d_flags + (N_COLS - 1), sizeof(int), cudaMemcpyDeviceToHost, st));
        HALO_CUDA(cudaStreamSynchronize(st));
        for (int i = 0; i < N_IN; ++i) if (!lane1 || i < COL_W || i >= COL_W + NW) h_flags[i] = f_all[i];
    }
    HALO_CUDA(cudaMemsetAsync(d_flags + (N_COLS - 1), 0, sizeof(int), st));
    const bool ids_derived = !ids_not_affine && !dense;
    lap("round 0 (upload, iNTTs)");

    Transcript<C> tr(LABEL_PLONK);

    // ---- round 1 (:114-115): the 16 witness commitments as one batched MSM over the resident generators
    if (lane1) {                                           // started beside round 0
        helper1.join();
        lane_finish(lane1);
        if (rc_w) { last_err() = err_w; return rc_w; }
    } else if (int e = commit_cols(col(COL_W), NW, [&](int i) { return nonzero(COL_W + i); }, o_Cws)) return e;
    for (int i = 0; i < NW; ++i) tr.absorb_point(H::from_words(o_Cws + 12 * i));
    lap("round 1 (16 commits)");

    // ---- round 3 (:129-162): grand product
    Challenges<S> ch;
    ch.beta = tr.challenge();
    ch.gamma = tr.challenge();
    {   // values on H of the live columns among w_0..7, id_0..7, sigma_0..7 (zero columns and derived ids are not transformed)
        ColMap pm;
        memset(&pm, 0, sizeof pm);
        int live3 = 0;
        const int src_col[3] = {COL_W, COL_ID, COL_SIG};
        for (int g3 = 0; g3 < 3; ++g3)
            for (int i = 0; i < NS; ++i) {
                const int c = src_col[g3] + i;
                const bool skip = !nonzero(c) || (ids_derived && g3 == 1 && i > 0);
                if (skip) { pm.slot[g3 * NS + i] = 255; continue; }
                HALO_CUDA(cudaMemcpyAsync(tmp + (size_t)live3 * n, col(c), n * 32, cudaMemcpyDeviceToDevice, st));
                pm.slot[g3 * NS + i] = (unsigned char)live3++;
            }
        for (int i = 0; i < 3 * NS; ++i) if (pm.slot[i] == 255) pm.slot[i] = (unsigned char)live3;      // the shared zero column
        if (int e = ntt_device<S>(tmp, ev, (int)lg_n, (size_t)live3, false, nullptr)) return e;
        if (live3 < 3 * NS) HALO_CUDA(cudaMemsetAsync(ev + (size_t)live3 * n, 0, n * 32, st));
        SF* num = f0; SF* den = f0 + n;
        plonk_perm_kernel<S><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(ev, n, ch.beta, ch.gamma, num, den,
                                                                        ids_derived ? fe_mul<S>(id_step, ch.beta) : fe_zero<S>(), pm);
    }
    SF* num = f0; SF* den = f0 + n;
    R.launches += 1;
    if (int e = scan_mul_device<S>(num, n, misc, false)) return e;
    if (int e = scan_mul_device<S>(den, n, misc, true)) return e;
    SF total;
    HALO_CUDA(cudaMemcpyAsync(&total, den, 32, cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    HALO_CHECK(!fe_is_zero<S>(total), HALO_EINVAL, "plonk: division by zero in the grand product");
    plonk_z_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(num, den, n, fe_inv<S>(total), tmp, tmp + n);
    R.launches += 1;
    HALO_CUDA(cudaGetLastError());
    if (int e = ntt_device<S>(tmp, col(COL_Z), (int)lg_n, 2, true, nullptr)) return e;          // z, z_omega
    plonk_nonzero_kernel<S><<<dim3(32, N_COLS - N_IN), 256, 0, st>>>(col(N_IN), n, d_flags + N_IN);
    R.launches += 1;
    HALO_CUDA(cudaMemcpyAsync(h_flags + N_IN, d_flags + N_IN, (N_COLS - N_IN) * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (int e = commit_run(col(COL_Z), 1, o_Cz)) return e;
    tr.absorb_point(H::from_words(o_Cz));
    lap("round 3 (z)");

    // ---- round 4 (:170-268): numerator on the 16 cosets of H, one interpolation, coefficient-form quotient
    ch.alpha = tr.challenge();
    ch.alpha2 = fe_sqr<S>(ch.alpha);
    for (int i = 0; i < 9; ++i) ch.M[i] = fe_from_words<S>(poseidon_data::POSEIDON_MDS[S][i]);
    unsigned gate_mask = 0;
    for (int i = 0; i < 5; ++i) if (nonzero(COL_Q + 5 + i)) gate_mask |= 1u << i;
    const SF w16n = root_of_unity<S>((int)lg_n + 4);
    // only the nonzero columns are staged and transformed (compactly, in column order); the zero ones share slot nnz
    ColMap map, list;
    int nnz = 0;
    memset(&map, 0, sizeof map); memset(&list, 0, sizeof list);
    // z_omega is read from z at the next point of the coset; id_1..7 come from id_0 when they differ by constants
    auto live = [&](int c) {
        if (c == COL_ZW) return false;
        if (ids_derived && c > COL_ID && c < COL_ID + NS) return false;
        return nonzero(c);
    };
    ch.id_step_beta = ids_derived ? fe_mul<S>(id_step, ch.beta) : fe_zero<S>();
    for (int c = 0; c < N_EVAL; ++c) if (live(c)) { list.slot[nnz] = (unsigned char)c; map.slot[c] = (unsigned char)nnz; ++nnz; }
    for (int c = 0; c < N_EVAL; ++c) if (!live(c)) map.slot[c] = (unsigned char)nnz;
    if (nnz < N_EVAL) HALO_CUDA(cudaMemsetAsync(ev + (size_t)nnz * n, 0, n * 32, st));
    auto numerator_on_coset = [&](unsigned c, SF* dst, unsigned slot, unsigned stride, const SF* scale) -> int {
        SF g = one;                                        // w16n^c
        for (unsigned i = 0; i < c; ++i) g = fe_mul<S>(g, w16n);
        if (c == 0) {
            for (int i = 0; i < nnz;) {                    // runs of adjacent source columns
                int j = i + 1;
                while (j < nnz && list.slot[j] == list.slot[j - 1] + 1) ++j;
                HALO_CUDA(cudaMemcpyAsync(tmp + (size_t)i * n, col(list.slot[i]), (size_t)(j - i) * n * 32, cudaMemcpyDeviceToDevice, st));
                i = j;
            }
        } else {
            SF g1024 = g;
            for (int i = 0; i < 10; ++i) g1024 = fe_sqr<S>(g1024);
            plonk_powers_kernel<S><<<1, 128, 0, st>>>(gtab, 1024, g);
            plonk_powers_kernel<S><<<(unsigned)((n / 1024 / 16 + 128) / 128), 128, 0, st>>>(gtab + 1024, (n + 1023) / 1024, g1024);
            plonk_coset_stage_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(cols, tmp, n, nnz, gtab, gtab + 1024, list);
            R.launches += 3;
        }
        if (int e = ntt_device<S>(tmp, ev, (int)lg_n, (size_t)nnz, false, nullptr)) return e;
        plonk_numerator_kernel<S><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(ev, n, ch, gate_mask, dst, slot, stride,
                                                                              scale ? *scale : one, scale ? 1 : 0, map);
        R.launches += 1;
        HALO_CUDA(cudaGetLastError());
        return HALO_OK;
    };
    // f vanishes on H exactly when X^n - 1 divides it (always, for a satisfying witness).  Then t = f / Z_H has
    // degree < 8n and is interpolated from the 8 odd cosets, which form the coset w16n * (8n domain) and on which
    // Z_H is the constant w16^c - 1.  Otherwise the reference still returns the polynomial quotient (protocol.rs:256
    // drops the remainder), which takes f on all 16 cosets and the coefficient-form division.
    if (residual_out) {
        // the gate constraints alone (alpha = 0 removes the permutation terms): n values f_GC(omega^j), j = 0 .. n-1
        ch.alpha = fe_zero<S>(); ch.alpha2 = fe_zero<S>();
        if (int e = numerator_on_coset(0, f1, 0, 1, nullptr)) return e;
        HALO_CUDA(cudaMemcpyAsync(residual_out, f1, n * 32, cudaMemcpyDeviceToHost, st));
        HALO_CUDA(cudaStreamSynchronize(st));
        if (trace) { cudaEventDestroy(ev0); cudaEventDestroy(ev1); }
        return HALO_OK;
    }
    if (int e = numerator_on_coset(0, f1, 0, 1, nullptr)) return e;
    HALO_CUDA(cudaMemsetAsync(d_flags, 0, sizeof(int), st));
    plonk_nonzero_kernel<S><<<dim3(64, 1), 256, 0, st>>>(f1, n, d_flags);
    R.launches += 1;
    int not_divisible = 0;
    HALO_CUDA(cudaMemcpyAsync(&not_divisible, d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    SF* tq = f0;                                           // 16 chunks of n coefficients (t_split, :509-517)
    if (!not_divisible && R.tune[5] != 1) {
        const SF w16 = root_of_unity<S>(4);
        SF zh = w16;                                       // w16^c for c = 1, 3, ...
        const SF w16sq = fe_sqr<S>(w16);
        const bool split = sharded && 8 % world == 0;      // rank r owns the cosets [8 r / world, 8 (r + 1) / world)
        for (unsigned k = 0; k < 8; ++k) {
            SF inv = fe_inv<S>(fe_sub<S>(zh, one));
            if (!split) { if (int e = numerator_on_coset(2 * k + 1, f1, k, 8, &inv)) return e; }
            else if ((int)(k * world / 8) == rank) { if (int e = numerator_on_coset(2 * k + 1, f0 + (size_t)k * n, 0, 1, &inv)) return e; }
            zh = fe_mul<S>(zh, w16sq);
        }
        if (split) {                                       // coset-major blocks, gathered in place, then interleaved for the 8n transform
            const size_t blk = (size_t)(8 / world) * n;
            if (int e = dist_allgather_dev(f0 + (size_t)rank * blk, f0, blk * 32)) return e;
            plonk_interleave_kernel<S><<<(unsigned)((8 * n + 255) / 256), 256, 0, st>>>(f0, n, 8, f1);
            R.launches += 1;
        }
        lap("round 4 (quotient on 8 cosets)");
        if (int e = ntt_device<S>(f1, tq, (int)lg_n + 3, 1, true, &w16n)) return e;
        HALO_CUDA(cudaMemsetAsync(tq + 8 * n, 0, 8 * n * 32, st));
    } else {
        const bool split = sharded && 16 % world == 0;
        for (unsigned c = 0; c < 16; ++c) {
            if (!split) { if (int e = numerator_on_coset(c, f0, c, 16, nullptr)) return e; }
            else if ((int)(c * world / 16) == rank) { if (int e = numerator_on_coset(c, f1 + (size_t)c * n, 0, 1, nullptr)) return e; }
        }
        if (split) {
            const size_t blk = (size_t)(16 / world) * n;
            if (int e = dist_allgather_dev(f1 + (size_t)rank * blk, f1, blk * 32)) return e;
            plonk_interleave_kernel<S><<<(unsigned)((16 * n + 255) / 256), 256, 0, st>>>(f1, n, 16, f0);
            R.launches += 1;
        }
        lap("round 4 (numerator, 16 cosets)");
        if (int e = ntt_device<S>(f0, f1, (int)lg_n + 4, 1, true, nullptr)) return e;
        plonk_quotient_kernel<S><<<(unsigned)((16 * n + 255) / 256), 256, 0, st>>>(f1, 16 * n, n, tq);
        R.launches += 1;
        HALO_CUDA(cudaGetLastError());
    }
    {   // on the 8-coset path deg t < 8n: chunks 8..15 are zero and commit to the identity
        // and a quotient of low degree (few gates in use) leaves more of them zero: flag the chunks like the columns
        const int nt_live = (!not_divisible && R.tune[5] != 1 && !dense) ? NT / 2 : NT;
        int t_flags[NT];
        HALO_CUDA(cudaMemsetAsync(d_flags, 0, NT * sizeof(int), st));
        plonk_nonzero_kernel<S><<<dim3(32, nt_live), 256, 0, st>>>(tq, n, d_flags);
        R.launches += 1;
        HALO_CUDA(cudaMemcpyAsync(t_flags, d_flags, nt_live * sizeof(int), cudaMemcpyDeviceToHost, st));
        HALO_CUDA(cudaStreamSynchronize(st));
        if (int e = commit_cols(tq, nt_live, [&](int i) { return dense || t_flags[i] != 0; }, o_Cts)) return e;
        const Jac<C> id = xyzz_to_jac<C>(xyzz_identity<C>());
        for (int i = nt_live; i < NT; ++i) memcpy(o_Cts + 12 * i, &id, 96);
    }
    for (int i = 0; i < NT; ++i) tr.absorb_point(H::from_words(o_Cts + 12 * i));
    lap("round 4 (quotient, 16 commits)");

    // ---- round 5 (:273-333)
    const SF zeta = tr.challenge();
    LinComb<S> lc;
    lc.count = 0;
    SF zp = one;
    auto push = [&](const SF* c) { lc.col[lc.count] = c; lc.coef[lc.count] = zp; ++lc.count; zp = fe_mul<S>(zp, zeta); };
    for (int i = 0; i < NQ; ++i) push(col(COL_Q + i));
    for (int i = 0; i < NW; ++i) push(col(COL_W + i));
    for (int i = 0; i < NT; ++i) push(tq + (size_t)i * n);
    push(col(COL_Z));
    SF* d_r = f1;                                          // r, r_omega: 2 x n
    plonk_lincomb_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(lc, n, d_r);
    lc.count = 0; zp = one;
    for (int i = 0; i < 3; ++i) push(col(COL_W + i));
    push(col(COL_Z));
    plonk_lincomb_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(lc, n, d_r + n);
    R.launches += 2;
    HALO_CUDA(cudaGetLastError());
    const SF xi = tr.challenge();
    const SF xi_omega = fe_mul<S>(xi, omega);

    // the 78 evaluations (:315-323) before the openings reuse the workspaces
    SF* pw = ev;                                           // powers of xi | powers of xi*omega
    plonk_powers_kernel<S><<<(unsigned)((n / 16 + 128) / 128), 128, 0, st>>>(pw, n, xi);
    plonk_powers_kernel<S><<<(unsigned)((n / 16 + 128) / 128), 128, 0, st>>>(pw + n, n, xi_omega);
    EvalSet<S> es;
    int ne = 0;
    auto ev_at = [&](const SF* c, bool shifted) { es.col[ne] = c; es.pw[ne] = shifted ? pw + n : pw; ++ne; };
    for (int i = 0; i < NW; ++i) ev_at(col(COL_W + i), false);
    for (int i = 0; i < NR; ++i) ev_at(col(COL_R + i), false);
    for (int i = 0; i < NQ; ++i) ev_at(col(COL_Q + i), false);
    for (int i = 0; i < NT; ++i) ev_at(tq + (size_t)i * n, false);
    for (int i = 0; i < NS; ++i) ev_at(col(COL_ID + i), false);
    for (int i = 0; i < NS; ++i) ev_at(col(COL_SIG + i), false);
    ev_at(col(COL_Z), false);
    ev_at(col(COL_Z), true);
    for (int i = 0; i < 3; ++i) ev_at(col(COL_WOM + i), false);
    SF* partial = misc + 2 * SCAN_SEG + 64;
    plonk_eval_kernel<S><<<dim3(EV_BLOCKS, ne), 256, 0, st>>>(es, n, partial);
    R.launches += 3;
    HALO_CUDA(cudaGetLastError());
    std::vector<SF> h_part((size_t)ne * EV_BLOCKS);
    HALO_CUDA(cudaMemcpyAsync(h_part.data(), partial, h_part.size() * 32, cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    for (int i = 0; i < ne; ++i) {
        SF acc = fe_zero<S>();
        for (int b = 0; b < EV_BLOCKS; ++b) acc = fe_add<S>(acc, h_part[(size_t)i * EV_BLOCKS + b]);
        H::put_scalar(acc, o_vs + 4 * i);
    }
    lap("round 5 (r, r_omega, 78 evals)");

    // Instance::open(r, xi) and Instance::open(r_omega, xi*omega) (:295-296): commit, then the IPA on resident data
    const unsigned iw = (unsigned)instance_words(lg_n);
    std::vector<uint64_t> insts(3 * (size_t)iw);
    memcpy(insts.data(), acc_prev, iw * 8);
    // The two openings are independent (different polynomials, different points, their own transcripts) and each is a
    // chain of lg n rounds whose last 15 are latency-bound and leave most of the GPU idle: the second one runs on a
    // helper thread in the library's second execution context (lane_get: own stream, workspaces, pinned buffer) beside
    // the first.  d_r is complete here (the evaluations above synchronised the stream).  Tuning key 13 = 1: one after
    // the other.
    auto open_k = [&](int k, bool shard_open = false) -> int {
        Runtime& Rk = rt();                                 // the main runtime, or the lane inside the helper thread
        uint64_t* q = insts.data() + (size_t)(1 + k) * iw;
        SF* cs;
        if (int e = ws_get(WS_IPA_C, n * 32, (void**)&cs)) return e;
        HALO_CUDA(cudaMemcpyAsync(cs, d_r + (size_t)k * n, n * 32, cudaMemcpyDeviceToDevice, Rk.stream));
        if (int e = msm_device<C>(Rk.pp[C].d_gs, cs, n, 1, q, 0)) return e;
        q[12] = d;
        const SF& at = k ? xi_omega : xi;
        H::put_scalar(at, q + 13);
        if (int e = pcdl_open<C>(nullptr, n, q, d, (const uint64_t*)&at, nullptr, nullptr, nullptr, nullptr, q + 21, q + 17, true, shard_open))
            return e;
        memcpy(k ? o_piro : o_pir, q + 21, evalproof_words(lg_n) * 8);
        return HALO_OK;
    };
    if (sharded && dist_half_available()) {
        // four or more GPUs: opening 0 sharded over the first HALF of the ranks, opening 1 over the second half, side by
        // side (halves of the communicator, ncclCommSplit); the two instances then travel over the full communicator
        const int hw = world / 2, h = rank / hw;
        if (int e = dist_half_enter()) return e;
        const int rc_h = open_k(h, true);
        dist_half_leave();
        if (rc_h) return rc_h;
        std::vector<uint64_t> mine(iw), all((size_t)iw * world);
        memcpy(mine.data(), insts.data() + (size_t)(1 + h) * iw, iw * 8);
        if (int e = dist_allgather_host(mine.data(), all.data(), (size_t)iw * 8)) return e;
        for (int k = 0; k < 2; ++k) {
            memcpy(insts.data() + (size_t)(1 + k) * iw, all.data() + (size_t)(k * hw) * iw, iw * 8);
            memcpy(k ? o_piro : o_pir, insts.data() + (size_t)(1 + k) * iw + 21, evalproof_words(lg_n) * 8);
        }
    } else if (sharded) {
        // opening k on rank k, the instances travel as one all-gather.  (Measured on 8 GPUs: an opening sharded over all
        // ranks takes 29.7 ms against 46 ms on one GPU -- its 15 tail rounds do not shard -- so two sharded openings one
        // after the other, 59 ms, lose against two unsharded ones side by side.)
        std::vector<uint64_t> mine(iw, 0), all((size_t)iw * world);
        if (rank < 2) {
            if (int e = open_k(rank)) return e;
            memcpy(mine.data(), insts.data() + (size_t)(1 + rank) * iw, iw * 8);
        }
        if (int e = dist_allgather_host(mine.data(), all.data(), (size_t)iw * 8)) return e;
        for (int k = 0; k < 2; ++k) {
            memcpy(insts.data() + (size_t)(1 + k) * iw, all.data() + (size_t)k * iw, iw * 8);
            memcpy(k ? o_piro : o_pir, insts.data() + (size_t)(1 + k) * iw + 21, evalproof_words(lg_n) * 8);
        }
    } else {
        Runtime* lane = nullptr;
        if (R.tune[13] != 1 && !R.is_lane && pcdl_open_prepare<C>(n) == HALO_OK) lane = lane_get();
        int rc1 = HALO_OK;
        std::string err1;
        Workers helper;
        if (lane) {
            helper.spawn(1, [&]() {
                LaneScope scope(lane);
                try {
                    rc1 = cudaSetDevice(lane->device) == cudaSuccess ? open_k(1) : HALO_ECUDA;
                } catch (...) { rc1 = HALO_ECUDA; last_err() = "plonk: exception in the second opening"; }
                if (rc1) err1 = last_err();
                cudaStreamSynchronize(lane->stream);
            });
        }
        const bool concurrent = helper.size() == 1;
        const int rc0 = open_k(0);
        helper.join();
        if (lane) lane_finish(lane);
        if (rc0) return rc0;
        if (concurrent) { if (rc1) { last_err() = err1; return rc1; } }
        else if (int e = open_k(1)) return e;
    }
    lap("round 5 (two openings)");
    if (int e = acc_prover<C>(insts.data(), 3, o_acc, sharded)) return e;   // sharded: its opening's upper rounds over all ranks
    lap("round 5 (accumulate)");
    if (trace) { cudaEventDestroy(ev0); cudaEventDestroy(ev1); }
    return HALO_OK;
}
template int plonk_prove<0>(unsigned, const uint64_t*, const uint64_t*, const uint64_t*, size_t, const uint64_t*, uint64_t*, const void*, bool, uint64_t*);
template int plonk_prove<1>(unsigned, const uint64_t*, const uint64_t*, const uint64_t*, size_t, const uint64_t*, uint64_t*, const void*, bool, uint64_t*);

// ---- circuit columns kept on the device across proofs: the analogue of the reference holding `x: PlonkCircuit` /
// the Trace's polynomials for every proof of the same circuit (an IVC chain proves one circuit over and over)
namespace {
struct ResidentCircuit { int curve; unsigned lg_n; void* d; };
std::vector<ResidentCircuit>& circuits() { static std::vector<ResidentCircuit> v; return v; }
}  // namespace

int plonk_circuit_load(int curve, unsigned lg_n, const uint64_t* polys, int* handle) {
    HALO_CHECK(lg_n >= 1 && lg_n <= 22, HALO_EINVAL, "plonk: rows must be a power of two in [2, 2^22]");
    const size_t n = (size_t)1 << lg_n;
    ResidentCircuit rc{curve, lg_n, nullptr};
    HALO_CUDA(cudaMalloc(&rc.d, (size_t)(N_IN - NW) * n * 32));
    char* d = (char*)rc.d;
    int e = upload_staged(d, polys, (size_t)NQ * n * 32);
    if (!e) e = upload_staged(d + (size_t)NQ * n * 32, polys + (size_t)COL_R * n * 4, (size_t)(NR + 2 * NS) * n * 32);
    if (!e && cudaStreamSynchronize(rt().stream) != cudaSuccess) e = HALO_ECUDA;
    if (e) { cudaFree(rc.d); return e; }
    auto& v = circuits();
    size_t slot = 0;
    while (slot < v.size() && v[slot].d) ++slot;
    if (slot == v.size()) v.push_back(rc); else v[slot] = rc;
    *handle = (int)slot + 1;
    return HALO_OK;
}

int plonk_circuit_free(int handle) {
    auto& v = circuits();
    HALO_CHECK(handle >= 1 && (size_t)handle <= v.size() && v[handle - 1].d, HALO_EINVAL, "plonk: unknown circuit handle");
    HALO_CUDA(cudaStreamSynchronize(rt().stream));
    HALO_CUDA(cudaFree(v[handle - 1].d));
    v[handle - 1].d = nullptr;
    return HALO_OK;
}

void plonk_circuits_release() {
    for (auto& c : circuits()) if (c.d) { cudaFree(c.d); c.d = nullptr; }
    circuits().clear();
}

int plonk_circuit_shape(int handle, int* curve, unsigned* lg_n) {
    auto& v = circuits();
    HALO_CHECK(handle >= 1 && (size_t)handle <= v.size() && v[handle - 1].d, HALO_EINVAL, "plonk: unknown circuit handle");
    *curve = v[handle - 1].curve; *lg_n = v[handle - 1].lg_n;
    return HALO_OK;
}
int plonk_prove_resident(int handle, const uint64_t* w_polys, const uint64_t* w_evals, const uint64_t* pub, size_t n_pub,
                         const uint64_t* acc_prev, uint64_t* out, bool shard) {
    auto& v = circuits();
    HALO_CHECK(handle >= 1 && (size_t)handle <= v.size() && v[handle - 1].d, HALO_EINVAL, "plonk: unknown circuit handle");
    const ResidentCircuit& rc = v[handle - 1];
    return rc.curve == 0 ? plonk_prove<0>(rc.lg_n, w_polys, w_evals, pub, n_pub, acc_prev, out, rc.d, shard, nullptr)
                         : plonk_prove<1>(rc.lg_n, w_polys, w_evals, pub, n_pub, acc_prev, out, rc.d, shard, nullptr);
}

namespace {
// the proof's claimed evaluations as the value source of the gate expressions
template <int F> struct ProofView {
    const Fe<F>*ws, *rs, *qs, *wom; Fe<F> pi_xi;
    HD Fe<F> q(int i) const { return qs[i]; }
    HD Fe<F> w(int i) const { return ws[i]; }
    HD Fe<F> r(int i) const { return rs[i]; }
    HD Fe<F> nw(int i) const { return wom[i]; }
    HD Fe<F> pi() const { return pi_xi; }
};
}  // namespace

// verify_succinct (protocol.rs:357-491) and verify (:493-502).  C_qs = circuit.Cs.qs (10 Jacobian points).
template <int C>
int plonk_verify(unsigned lg_n, const uint64_t* C_qs, const uint64_t* pub, size_t n_pub, size_t public_input_count,
                 const uint64_t* acc_prev, const uint64_t* proof, int succinct_only) {
    typedef Host<C> H;
    typedef typename H::SF SF;
    typedef typename H::Pt Pt;
    constexpr int S = H::S;
    HALO_CHECK(lg_n >= 1 && lg_n <= 22, HALO_EINVAL, "plonk: rows must be a power of two in [2, 2^22]");
    const size_t rows = (size_t)1 << lg_n, d = rows - 1;
    const SF one = fe_one<S>();
    const SF nS = fe_from_u64<S>((u64)rows);
    HALO_CHECK(n_pub == public_input_count, HALO_EREJECT, "public_inputs.len() == circuit.public_input_count");
    const SF* vs = (const SF*)proof;
    const SF *v_ws = vs, *v_rs = vs + 16, *v_qs = vs + 31, *v_ts = vs + 41, *v_ids = vs + 57, *v_sig = vs + 65;
    const SF v_z = vs[73], v_zw = vs[74];
    const SF* v_wom = vs + 75;
    for (int i = 0; i < 78; ++i) HALO_CHECK(H::scalar_is_valid(vs[i]), HALO_EINVAL, "plonk: evaluation not below the modulus");
    const uint64_t* o_Cws = proof + 78 * 4;
    const uint64_t* o_Cts = o_Cws + NW * 12;
    const uint64_t* o_Cz = o_Cts + NT * 12;
    const uint64_t* o_pir = o_Cz + 12;
    const uint64_t* o_piro = o_pir + evalproof_words(lg_n);
    const uint64_t* o_acc = o_piro + evalproof_words(lg_n);
    // the reference's types guarantee canonical field elements and on-curve points; raw words have to be checked
    // (the two opening proofs and both accumulators are validated as instances inside acc_verifier)
    for (int i = 0; i < NW + NT + 1; ++i)
        HALO_CHECK(H::point_words_valid(o_Cws + 12 * i), HALO_EREJECT, "plonk: a commitment of the proof is not a point of the curve");
    for (int i = 0; i < NQ; ++i)
        HALO_CHECK(H::point_words_valid(C_qs + 12 * i), HALO_EINVAL, "plonk: a circuit commitment is not a point of the curve");
    for (size_t j = 0; j < n_pub; ++j)
        HALO_CHECK(H::scalar_words_valid(pub + 4 * j), HALO_EINVAL, "plonk: public input not below the modulus");

    Transcript<C> tr(LABEL_PLONK);
    Pt Cws[NW], Cts[NT], Cz = H::from_words(o_Cz);
    for (int i = 0; i < NW; ++i) { Cws[i] = H::from_words(o_Cws + 12 * i); tr.absorb_point(Cws[i]); }
    const SF beta = tr.challenge(), gamma = tr.challenge();
    tr.absorb_point(Cz);
    const SF alpha = tr.challenge();
    for (int i = 0; i < NT; ++i) { Cts[i] = H::from_words(o_Cts + 12 * i); tr.absorb_point(Cts[i]); }
    const SF zeta = tr.challenge(), xi = tr.challenge();
    SF xi_n = xi;
    for (unsigned i = 0; i < lg_n; ++i) xi_n = fe_sqr<S>(xi_n);
    const SF omega = root_of_unity<S>((int)lg_n);
    const SF xi_omega = fe_mul<S>(xi, omega);

    SF f_prime = one, g_prime = one;
    for (int i = 0; i < NS; ++i) {
        SF wg = fe_add<S>(v_ws[i], gamma);
        f_prime = fe_mul<S>(f_prime, fe_add<S>(wg, fe_mul<S>(beta, v_ids[i])));
        g_prime = fe_mul<S>(g_prime, fe_add<S>(wg, fe_mul<S>(beta, v_sig[i])));
    }
    // public_input_eval_generic (:565-589)
    const SF zh = fe_sub<S>(xi_n, one);
    SF pi_xi = fe_zero<S>(), omega_j = omega;
    for (size_t j = 0; j < n_pub; ++j) {
        SF den = fe_mul<S>(nS, fe_sub<S>(xi, omega_j));
        HALO_CHECK(!fe_is_zero<S>(den), HALO_EINVAL, "plonk: division by zero in the public-input evaluation");
        SF lj = fe_mul<S>(fe_mul<S>(zh, omega_j), fe_inv<S>(den));
        pi_xi = fe_sub<S>(pi_xi, fe_mul<S>(lj, H::scalar(pub + 4 * j)));
        omega_j = fe_mul<S>(omega_j, omega);
    }
    SF M[9];
    for (int i = 0; i < 9; ++i) M[i] = fe_from_words<S>(poseidon_data::POSEIDON_MDS[S][i]);
    ProofView<S> pv{v_ws, v_rs, v_qs, v_wom, pi_xi};
    const SF f_gc = f_gc_point<S, ProofView<S>>(pv, M, 31u);
    SF den = fe_mul<S>(nS, fe_sub<S>(xi, omega));
    HALO_CHECK(!fe_is_zero<S>(den), HALO_EINVAL, "plonk: division by zero in l1");
    const SF l1 = fe_mul<S>(fe_mul<S>(omega, zh), fe_inv<S>(den));
    const SF f_cc1 = fe_mul<S>(l1, fe_sub<S>(v_z, one));
    const SF f_cc2 = fe_sub<S>(fe_mul<S>(v_z, f_prime), fe_mul<S>(v_zw, g_prime));
    const SF f = fe_add<S>(f_gc, fe_add<S>(fe_mul<S>(alpha, f_cc1), fe_mul<S>(fe_sqr<S>(alpha), f_cc2)));
    SF t = v_ts[0], acc = xi_n;
    for (int i = 1; i < NT; ++i) { t = fe_add<S>(t, fe_mul<S>(acc, v_ts[i])); acc = fe_mul<S>(acc, xi_n); }
    HALO_CHECK(fe_eq<S>(f, fe_mul<S>(t, zh)), HALO_EREJECT, "T(xi) != (F_GC(xi) + a F_CC1(xi) + a^2 F_CC2(xi)) / Z_H(xi)");

    // the two batched openings (:455-477): v_r, C_r over (qs, ws, ts, z) and v_r_omega, C_r_omega over (w0..2, z)
    SF v_r = v_qs[0];
    Pt C_r = H::from_words(C_qs);
    SF a = zeta;
    auto fold = [&](const SF& v, const Pt& P) { v_r = fe_add<S>(v_r, fe_mul<S>(v, a)); C_r = H::add(C_r, H::mul(P, a)); a = fe_mul<S>(a, zeta); };
    for (int i = 1; i < NQ; ++i) fold(v_qs[i], H::from_words(C_qs + 12 * i));
    for (int i = 0; i < NW; ++i) fold(v_ws[i], Cws[i]);
    for (int i = 0; i < NT; ++i) fold(v_ts[i], Cts[i]);
    fold(v_z, Cz);
    SF v_ro = v_wom[0];
    Pt C_ro = Cws[0];
    a = zeta;
    for (int i = 1; i < 3; ++i) { v_ro = fe_add<S>(v_ro, fe_mul<S>(v_wom[i], a)); C_ro = H::add(C_ro, H::mul(Cws[i], a)); a = fe_mul<S>(a, zeta); }
    v_ro = fe_add<S>(v_ro, fe_mul<S>(v_zw, a));
    C_ro = H::add(C_ro, H::mul(Cz, a));

    const unsigned iw = (unsigned)instance_words(lg_n);
    HALO_CHECK(acc_prev[12] == d && o_acc[12] == d, HALO_EREJECT, "d_i != d");
    std::vector<uint64_t> insts(3 * (size_t)iw);
    memcpy(insts.data(), acc_prev, iw * 8);
    for (int k = 0; k < 2; ++k) {
        uint64_t* q = insts.data() + (size_t)(1 + k) * iw;
        H::to_words(k ? C_ro : C_r, q);
        q[12] = d;
        H::put_scalar(k ? xi_omega : xi, q + 13);
        H::put_scalar(k ? v_ro : v_r, q + 17);
        memcpy(q + 21, k ? o_piro : o_pir, evalproof_words(lg_n) * 8);
    }
    if (int e = acc_verifier<C>(insts.data(), 3, o_acc)) return e;
    if (succinct_only) return HALO_OK;
    return pcdl_check<C>(o_acc);                           // acc::decider (acc.rs:215-217)
}
template int plonk_verify<0>(unsigned, const uint64_t*, const uint64_t*, size_t, size_t, const uint64_t*, const uint64_t*, int);
template int plonk_verify<1>(unsigned, const uint64_t*, const uint64_t*, size_t, size_t, const uint64_t*, const uint64_t*, int);

}  // namespace halo
