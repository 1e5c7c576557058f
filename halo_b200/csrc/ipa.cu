// PC_DL (inner-product-argument polynomial commitment) and AS_DL (its accumulation scheme) on sm_100a.
//
// Stands in for
//   crates/accumulation/src/pcdl.rs:266-287   commit
//   crates/accumulation/src/pcdl.rs:326-473   open_without_eval / open        (the lg n fold rounds run on the GPU)
//   crates/accumulation/src/pcdl.rs:483-554   succinct_check                  (O(lg n) host work)
//   crates/accumulation/src/pcdl.rs:563-583   check                           (h(X) expanded + committed on the GPU)
//   crates/accumulation/src/pcdl.rs:191-233   HPoly::{get_poly, eval}
//   crates/accumulation/src/acc.rs:71-110,128-217   common_subroutine / prover / verifier / decider
//   crates/group/src/group.rs:43-66           scalar_dot / construct_powers (device kernels below)
// Division of labour: everything whose size is n (generators, coefficient and power vectors, h(X)) lives in HBM and
// is touched only by kernels; the Fiat-Shamir transcript and the O(lg n) point arithmetic of the verifier stay on
// the host, which receives two points per round and sends one challenge back.
#include "runtime.cuh"
#include "transcript.cuh"
#include "host.cuh"
#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <thread>
#include <chrono>
#include <cstdio>
#include <vector>

extern "C" int halo_msm_proj(int, const uint64_t*, const uint64_t*, size_t, uint64_t*);

namespace halo {

// dist.cu (declared in host.cuh): the library's NCCL layer (one process per GPU)
void dist_range(size_t total, size_t* lo, size_t* hi);

template <int F> __device__ __forceinline__ Fe<F> g_ld(const Fe<F>* p) {
    Fe<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int F> __device__ __forceinline__ void g_st(Fe<F>* p, const Fe<F>& v) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    q[1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}


// ---------------------------------------------------------------------------------------------------------
// construct_powers (group.rs:58-66): out[j] = c0 * z^j, 16 consecutive powers per thread
// ---------------------------------------------------------------------------------------------------------
template <int F>
__global__ void __launch_bounds__(128) ipa_powers_kernel(Fe<F>* __restrict__ out, size_t n, Fe<F> z, Fe<F> c0) {
    size_t lo = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 16;
    if (lo >= n) return;
    u32 e[8] = {(u32)lo, (u32)(lo >> 32), 0, 0, 0, 0, 0, 0};
    Fe<F> w = fe_mul<F>(fe_pow<F>(z, e, 40), c0);
    size_t hi = lo + 16 < n ? lo + 16 : n;
    for (size_t k = lo; k < hi; ++k) { g_st(out + k, w); w = fe_mul<F>(w, z); }
}

// ---------------------------------------------------------------------------------------------------------
// scalar_dot (group.rs:43-45), both dot products of a round in one pass:
//   part[2*blk] = sum_j a_hi[j]*b_lo[j]  (<r(c), l(z)>),  part[2*blk+1] = sum_j a_lo[j]*b_hi[j]  (<l(c), r(z)>)
// With b == nullptr computes plain sums (used for reductions).  The host adds the per-block partials.
// ---------------------------------------------------------------------------------------------------------
template <int F>
__global__ void __launch_bounds__(256) ipa_dots_kernel(const Fe<F>* __restrict__ a, const Fe<F>* __restrict__ b, size_t m,
                                                       Fe<F>* __restrict__ part) {
    __shared__ Fe<F> sh[2][256];
    Fe<F> s0 = fe_zero<F>(), s1 = fe_zero<F>();
    for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < m; j += (size_t)gridDim.x * blockDim.x) {
        s0 = fe_add<F>(s0, fe_mul<F>(g_ld(a + m + j), g_ld(b + j)));
        s1 = fe_add<F>(s1, fe_mul<F>(g_ld(a + j), g_ld(b + m + j)));
    }
    sh[0][threadIdx.x] = s0; sh[1][threadIdx.x] = s1;
    __syncthreads();
    for (int d = 128; d >= 1; d >>= 1) {
        if ((int)threadIdx.x < d) {
            sh[0][threadIdx.x] = fe_add<F>(sh[0][threadIdx.x], sh[0][threadIdx.x + d]);
            sh[1][threadIdx.x] = fe_add<F>(sh[1][threadIdx.x], sh[1][threadIdx.x + d]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { g_st(part + 2 * blockIdx.x, sh[0][0]); g_st(part + 2 * blockIdx.x + 1, sh[1][0]); }
}
// single dot product <a, b> over n elements -> per-block partials
template <int F>
__global__ void __launch_bounds__(256) ipa_dot1_kernel(const Fe<F>* __restrict__ a, const Fe<F>* __restrict__ b, size_t n,
                                                       Fe<F>* __restrict__ part) {
    __shared__ Fe<F> sh[256];
    Fe<F> s = fe_zero<F>();
    for (size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x; j < n; j += (size_t)gridDim.x * blockDim.x)
        s = fe_add<F>(s, fe_mul<F>(g_ld(a + j), g_ld(b + j)));
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int d = 128; d >= 1; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = fe_add<F>(sh[threadIdx.x], sh[threadIdx.x + d]);
        __syncthreads();
    }
    if (threadIdx.x == 0) g_st(part + blockIdx.x, sh[0]);
}

// ---------------------------------------------------------------------------------------------------------
// one fold round (pcdl.rs:427-435), all three vectors in one kernel:
//   gs[j] = (gs[j] + xi * gs[m+j]).into_affine()      (uniform double-and-add: every lane has the same xi)
//   cs[j] += xi^-1 * cs[m+j]        zs[j] += xi * zs[m+j]
// ---------------------------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(128, 4) ipa_fold_kernel(Affine<C>* __restrict__ gs, Fe<CurveTraits<C>::SCALAR>* __restrict__ cs,
                                                          Fe<CurveTraits<C>::SCALAR>* __restrict__ zs, size_t m,
                                                          Bits256 xi_bits, Fe<CurveTraits<C>::SCALAR> xi,
                                                          Fe<CurveTraits<C>::SCALAR> xi_inv) {
    constexpr int S = CurveTraits<C>::SCALAR;
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= m) return;
    g_st(cs + j, fe_add<S>(g_ld(cs + j), fe_mul<S>(g_ld(cs + m + j), xi_inv)));
    g_st(zs + j, fe_add<S>(g_ld(zs + j), fe_mul<S>(g_ld(zs + m + j), xi)));
    Affine<C> P;
    P.x = g_ld(&gs[m + j].x); P.y = g_ld(&gs[m + j].y);
    XYZZ<C> acc = xyzz_identity<C>();
    for (int b = 254; b >= 0; --b) {
        acc = xyzz_dbl<C>(acc);
        if ((xi_bits.w[b >> 5] >> (b & 31)) & 1u) xyzz_madd<C>(acc, P, false);
    }
    Affine<C> Q;
    Q.x = g_ld(&gs[j].x); Q.y = g_ld(&gs[j].y);
    xyzz_madd<C>(acc, Q, false);
    Affine<C> r = xyzz_to_affine<C>(acc);
    g_st(&gs[j].x, r.x); g_st(&gs[j].y, r.y);
}

// ---------------------------------------------------------------------------------------------------------
// The same fold with the curve endomorphism: both Pasta curves have j = 0, so phi(x, y) = (beta*x, y) = lambda*(x, y)
// and xi = k1 + k2*lambda with |k1|, |k2| < 2^128 (host: glv_decompose).  xi*P = k1*P + k2*phi(P) needs half the
// doublings of the plain ladder.  The result stays in XYZZ form; xyzz_batch_affine_kernel normalises a run of
// points per thread with ONE inversion (Montgomery's trick) instead of one Fermat inversion per point.
// ---------------------------------------------------------------------------------------------------------
struct GlvScalar {
    u32 k1[5], k2[5];      // magnitudes, little-endian limbs
    u32 p1[5], n1[5];      // non-adjacent form of |k1|: digit +1 / -1 bit masks (a third of the digits are non-zero)
    u32 p2[5], n2[5];      // ... of |k2|
    int neg1, neg2;        // signs
    int nbits;             // number of NAF digits to scan
};
template <int C>
__global__ void __launch_bounds__(128, 4) ipa_fold_glv_kernel(const Affine<C>* __restrict__ gs,
                                                              Fe<CurveTraits<C>::SCALAR>* __restrict__ cs,
                                                              Fe<CurveTraits<C>::SCALAR>* __restrict__ zs, size_t m, GlvScalar g,
                                                              Fe<CurveTraits<C>::BASE> beta, Fe<CurveTraits<C>::SCALAR> xi,
                                                              Fe<CurveTraits<C>::SCALAR> xi_inv, XYZZ<C>* __restrict__ out) {
    constexpr int S = CurveTraits<C>::SCALAR, B = CurveTraits<C>::BASE;
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= m) return;
    g_st(cs + j, fe_add<S>(g_ld(cs + j), fe_mul<S>(g_ld(cs + m + j), xi_inv)));
    g_st(zs + j, fe_add<S>(g_ld(zs + j), fe_mul<S>(g_ld(zs + m + j), xi)));
    Affine<C> P1, P2;
    P1.x = g_ld(&gs[m + j].x); P1.y = g_ld(&gs[m + j].y);
    P2.x = fe_mul<B>(P1.x, beta);                         // phi(P); the identity (0,0) maps to itself
    P2.y = g.neg2 ? fe_neg<B>(P1.y) : P1.y;
    if (g.neg1) P1.y = fe_neg<B>(P1.y);
    if (fe_is_zero<B>(P1.x) && fe_is_zero<B>(P2.y)) { P1.y = fe_zero<B>(); P2.y = fe_zero<B>(); }   // keep (0,0) exact
    XYZZ<C> acc = xyzz_identity<C>();
    for (int b = g.nbits - 1; b >= 0; --b) {
        acc = xyzz_dbl<C>(acc);
        const u32 w = (u32)b >> 5, sh = (u32)b & 31;
        const u32 d1 = ((g.p1[w] | g.n1[w]) >> sh) & 1u, d2 = ((g.p2[w] | g.n2[w]) >> sh) & 1u;
        if (d1) xyzz_madd<C>(acc, P1, ((g.n1[w] >> sh) & 1u) != 0);
        if (d2) xyzz_madd<C>(acc, P2, ((g.n2[w] >> sh) & 1u) != 0);
    }
    Affine<C> Q;
    Q.x = g_ld(&gs[j].x); Q.y = g_ld(&gs[j].y);
    xyzz_madd<C>(acc, Q, false);
    XYZZ<C>* o = out + j;
    g_st(&o->X, acc.X); g_st(&o->Y, acc.Y); g_st(&o->ZZ, acc.ZZ); g_st(&o->ZZZ, acc.ZZZ);
}

// Two folds at once.  Delaying a fold by one round (the round in between computes its L and R over the unfolded
// generators with weights, see the tail kernels below) turns two ladders into one JOINT ladder:
//     g'[i] = B[i] + w1*B[i+L] + w2*B[i+2L] + w3*B[i+3L],   (w1, w2, w3) = (xi_b, xi_a, xi_a*xi_b)
// shares its 128 doublings between the three scalars and produces n/4 instead of n/2 + n/4 points: ~40 % fewer
// field multiplications per input generator than two radix-2 folds.  NS = 1 is the single delayed fold.
struct JointScalars { GlvScalar s[3]; int nbits; };
// The 3*NS operand coordinates (x, y, beta*x of each input) live in shared memory, one 16-byte lane per thread and
// half-element, so that the ladder itself fits 128 registers and four CTAs share an SM (16 warps instead of 12).
template <int F> __device__ __forceinline__ void s_put(uint4* sm, int slot, const Fe<F>& v) {
    sm[(2 * slot) * 128 + threadIdx.x] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    sm[(2 * slot + 1) * 128 + threadIdx.x] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}
template <int F> __device__ __forceinline__ Fe<F> s_get(const uint4* sm, int slot) {
    const uint4 a = sm[(2 * slot) * 128 + threadIdx.x], b = sm[(2 * slot + 1) * 128 + threadIdx.x];
    Fe<F> r;
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w; r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int C, int NS>
__global__ void __launch_bounds__(128, 4) ipa_materialize_kernel(const Affine<C>* __restrict__ Bs, size_t L, JointScalars js,
                                                                 Fe<CurveTraits<C>::BASE> beta, XYZZ<C>* __restrict__ out) {
    constexpr int B = CurveTraits<C>::BASE;
    extern __shared__ uint4 joint_sm[];                  // [3*NS slots][2 halves][128 threads]
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= L) return;
#pragma unroll
    for (int t = 0; t < NS; ++t) {
        const Fe<B> x = g_ld(&Bs[i + (size_t)(t + 1) * L].x), y = g_ld(&Bs[i + (size_t)(t + 1) * L].y);
        s_put<B>(joint_sm, 3 * t, x); s_put<B>(joint_sm, 3 * t + 1, y); s_put<B>(joint_sm, 3 * t + 2, fe_mul<B>(x, beta));
    }
    // ONE addition site in the loop body: with the six digit streams unrolled the body is ~70 inlined field
    // multiplications (> 300 KB of code) and the ladder runs out of the instruction cache
    XYZZ<C> acc = xyzz_identity<C>();
    for (int b = js.nbits - 1; b >= 0; --b) {
        acc = xyzz_dbl<C>(acc);
        const u32 w = (u32)b >> 5, sh = (u32)b & 31;
#pragma unroll 1
        for (int k = 0; k < 2 * NS; ++k) {
            const GlvScalar& g = js.s[k >> 1];
            const bool second = (k & 1) != 0;            // the phi(P) half: (beta*x, y); (0,0) stays the identity
            const u32 pos = ((second ? g.p2[w] : g.p1[w]) >> sh) & 1u, neg = ((second ? g.n2[w] : g.n1[w]) >> sh) & 1u;
            if (pos | neg) {
                Affine<C> Q;
                Q.x = s_get<B>(joint_sm, 3 * (k >> 1) + (second ? 2 : 0)); Q.y = s_get<B>(joint_sm, 3 * (k >> 1) + 1);
                xyzz_madd<C>(acc, Q, (neg != 0) != ((second ? g.neg2 : g.neg1) != 0));
            }
        }
    }
    Affine<C> Q0;
    Q0.x = g_ld(&Bs[i].x); Q0.y = g_ld(&Bs[i].y);
    xyzz_madd<C>(acc, Q0, false);
    XYZZ<C>* o = out + i;
    g_st(&o->X, acc.X); g_st(&o->Y, acc.Y); g_st(&o->ZZ, acc.ZZ); g_st(&o->ZZZ, acc.ZZZ);
}

// ---------------------------------------------------------------------------------------------------------
// The FIRST fold of an opening multiplies the upper half of pp.Gs -- fixed points -- by a fresh challenge.  With
// the table  tab[k][j] = 2^k * Gs[n/2 + j]  (affine, k-major so that a warp reads 32 consecutive points) the product
// xi * Gs[n/2 + j] = sum over the non-zero NAF digits of k1 and k2 of +-tab[k][j] (phi applied on the fly for k2)
// needs NO doublings: ~86 mixed additions instead of 128 doublings + ~86 additions.  Because the doublings are gone
// there is nothing to share between scalars, so this fold is taken at once (radix 2) and the radix-4 pairing starts
// one round later.  Width-3 NAF digits (+-1, +-3; a quarter of the positions non-zero) take the triples 3 * 2^k * P
// from a second set of rows: ~64 additions.  The table (2 * FOLD_ROWS * n/2 points, 9.1 GB at n = 2^20) is built
// once per parameter set.
// ---------------------------------------------------------------------------------------------------------
constexpr int FOLD_ROWS = 136, FOLD_SLICE = 34;
template <int C>
__global__ void __launch_bounds__(128) ipa_fold_table_rows_kernel(const Affine<C>* __restrict__ hi, XYZZ<C>* __restrict__ state,
                                                                  size_t m, int first, int rows, XYZZ<C>* __restrict__ out,
                                                                  XYZZ<C>* __restrict__ out3) {
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= m) return;
    XYZZ<C> P;
    if (first) { Affine<C> a; a.x = g_ld(&hi[j].x); a.y = g_ld(&hi[j].y); P = xyzz_from_affine<C>(a); }
    else { P.X = g_ld(&state[j].X); P.Y = g_ld(&state[j].Y); P.ZZ = g_ld(&state[j].ZZ); P.ZZZ = g_ld(&state[j].ZZZ); }
    for (int k = 0; k < rows; ++k) {
        XYZZ<C>* o = out + (size_t)k * m + j;
        g_st(&o->X, P.X); g_st(&o->Y, P.Y); g_st(&o->ZZ, P.ZZ); g_st(&o->ZZZ, P.ZZZ);
        XYZZ<C> D = xyzz_dbl<C>(P);
        xyzz_add<C>(P, D);                                 // 3 * 2^k * G
        o = out3 + (size_t)k * m + j;
        g_st(&o->X, P.X); g_st(&o->Y, P.Y); g_st(&o->ZZ, P.ZZ); g_st(&o->ZZZ, P.ZZZ);
        P = D;
    }
    g_st(&state[j].X, P.X); g_st(&state[j].Y, P.Y); g_st(&state[j].ZZ, P.ZZ); g_st(&state[j].ZZZ, P.ZZZ);
}
// out[j] = lo[j] + xi * hi[j] from the table of hi.  The host turns the two half-scalars into a list of additions
// (width-3 NAF: digits +-1 and +-3, a quarter of the positions non-zero): table row (k, or FOLD_ROWS + k for the
// triples) | phi << 9 | negate << 10.  One addition site; the next operand is loaded while the current addition runs.
struct FoldOps { unsigned short op[160]; int n; };
template <int C>
__global__ void __launch_bounds__(128, 4) ipa_fold_table_kernel(const Affine<C>* __restrict__ lo, const Affine<C>* __restrict__ tab,
                                                                size_t m, const __grid_constant__ FoldOps ops,
                                                                Fe<CurveTraits<C>::BASE> beta, XYZZ<C>* __restrict__ out) {
    constexpr int B = CurveTraits<C>::BASE;
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= m) return;
    Affine<C> Q0;
    Q0.x = g_ld(&lo[j].x); Q0.y = g_ld(&lo[j].y);
    XYZZ<C> acc = xyzz_from_affine<C>(Q0);
    Affine<C> Q;
    if (ops.n) { const Affine<C>* t = tab + (size_t)(ops.op[0] & 0x1ffu) * m + j; Q.x = g_ld(&t->x); Q.y = g_ld(&t->y); }
#pragma unroll 1
    for (int i = 0; i < ops.n; ++i) {
        const u32 op = ops.op[i];
        Affine<C> Qn;
        if (i + 1 < ops.n) { const Affine<C>* t = tab + (size_t)(ops.op[i + 1] & 0x1ffu) * m + j; Qn.x = g_ld(&t->x); Qn.y = g_ld(&t->y); }
        if (op & 0x200u) Q.x = fe_mul<B>(Q.x, beta);          // phi; (0,0) stays the identity
        xyzz_madd<C>(acc, Q, (op & 0x400u) != 0);
        Q = Qn;
    }
    XYZZ<C>* o = out + j;
    g_st(&o->X, acc.X); g_st(&o->Y, acc.Y); g_st(&o->ZZ, acc.ZZ); g_st(&o->ZZZ, acc.ZZZ);
}

// XYZZ -> affine for `m` points, BATCH consecutive points per thread sharing one inversion; identity -> (0,0)
template <int C, int BATCH>
__global__ void __launch_bounds__(128) xyzz_batch_affine_kernel(const XYZZ<C>* __restrict__ in, size_t m, Affine<C>* __restrict__ out) {
    constexpr int B = CurveTraits<C>::BASE;
    const size_t lo = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * BATCH;
    if (lo >= m) return;
    Fe<B> pre[BATCH];
    Fe<B> run = fe_one<B>();
#pragma unroll
    for (int i = 0; i < BATCH; ++i) {
        if (lo + i < m) {
            Fe<B> zz = g_ld(&in[lo + i].ZZ), zzz = g_ld(&in[lo + i].ZZZ);
            if (!fe_is_zero<B>(zz)) run = fe_mul<B>(run, fe_mul<B>(zz, zzz));
        }
        pre[i] = run;
    }
    Fe<B> inv = fe_inv<B>(run);
#pragma unroll
    for (int i = BATCH - 1; i >= 0; --i) {
        if (lo + i < m) {
            Fe<B> zz = g_ld(&in[lo + i].ZZ), zzz = g_ld(&in[lo + i].ZZZ);
            Affine<C> a;
            if (fe_is_zero<B>(zz)) { a.x = fe_zero<B>(); a.y = fe_zero<B>(); }
            else {
                Fe<B> dinv = i > 0 ? fe_mul<B>(inv, pre[i - 1]) : inv;       // 1 / (ZZ*ZZZ)
                inv = fe_mul<B>(inv, fe_mul<B>(zz, zzz));
                a.x = fe_mul<B>(g_ld(&in[lo + i].X), fe_mul<B>(dinv, zzz));
                a.y = fe_mul<B>(g_ld(&in[lo + i].Y), fe_mul<B>(dinv, zz));
            }
            g_st(&out[lo + i].x, a.x); g_st(&out[lo + i].y, a.y);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Fold-free tail.  Once the vectors are short, a round's fold is pure latency (a 128-step ladder per generator,
// ~1.1 ms however few there are).  From then on the generators are never folded again: with B the T generators
// at the switch and w the weights accumulated since (w = [1] at the switch),
//     g[i] = sum_u w[u] * B[i + u*2m]          (2m = current vector length, u < T/2m)
// so L = <r(c), l(g)> and R = <l(c), r(g)> are MSMs over the SAME T bases with complementary scalar vectors, one
// two-MSM pipeline per round; a fold only updates w (w'[2u] = w[u], w'[2u+1] = xi*w[u]); U = <w, B> at the end.
// The results are the same group elements as the reference's literal fold (pcdl.rs:427-429).
// ---------------------------------------------------------------------------------------------------------
template <int F>
__global__ void __launch_bounds__(256) ipa_tail_scalars_kernel(const Fe<F>* __restrict__ cs, const Fe<F>* __restrict__ w,
                                                               size_t m, size_t T, Fe<F>* __restrict__ sR, Fe<F>* __restrict__ sL) {
    size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (idx >= T) return;
    const size_t u = idx / (2 * m), i = idx - u * 2 * m;
    const Fe<F> wu = g_ld(w + u);
    if (i < m) { g_st(sL + idx, fe_mul<F>(g_ld(cs + m + i), wu)); g_st(sR + idx, fe_zero<F>()); }
    else { g_st(sR + idx, fe_mul<F>(g_ld(cs + i - m), wu)); g_st(sL + idx, fe_zero<F>()); }
}
template <int F>
__global__ void __launch_bounds__(256) ipa_tail_fold_kernel(Fe<F>* __restrict__ cs, Fe<F>* __restrict__ zs, size_t m,
                                                            const Fe<F>* __restrict__ w_in, size_t nu, Fe<F>* __restrict__ w_out,
                                                            Fe<F> xi, Fe<F> xi_inv) {
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j < m) {
        g_st(cs + j, fe_add<F>(g_ld(cs + j), fe_mul<F>(g_ld(cs + m + j), xi_inv)));
        g_st(zs + j, fe_add<F>(g_ld(zs + j), fe_mul<F>(g_ld(zs + m + j), xi)));
    }
    if (j < nu) { Fe<F> v = g_ld(w_in + j); g_st(w_out + 2 * j, v); g_st(w_out + 2 * j + 1, fe_mul<F>(v, xi)); }
}

// ---------------------------------------------------------------------------------------------------------
// hiding helpers (pcdl.rs:344-366):  p_bar = q * (X - z)  ->  pbar[i] = q[i-1] - z*q[i]  (q has d = n-1 coefficients)
// and  cs = p + a * p_bar
// ---------------------------------------------------------------------------------------------------------
template <int F>
__global__ void __launch_bounds__(256) ipa_pbar_kernel(const Fe<F>* __restrict__ q, size_t n, Fe<F> z, Fe<F>* __restrict__ pbar) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fe<F> v = i >= 1 ? g_ld(q + i - 1) : fe_zero<F>();
    if (i + 1 < n) v = fe_sub<F>(v, fe_mul<F>(g_ld(q + i), z));
    g_st(pbar + i, v);
}
template <int F>
__global__ void __launch_bounds__(256) ipa_axpy_kernel(Fe<F>* __restrict__ y, const Fe<F>* __restrict__ x, size_t n, Fe<F> a) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    g_st(y + i, fe_add<F>(g_ld(y + i), fe_mul<F>(g_ld(x + i), a)));
}

// ---------------------------------------------------------------------------------------------------------
// h(X) = prod_{i<lg n} (1 + xi_{lg n - i} X^{2^i})  (pcdl.rs:198-219): coefficient j is the product of the
// challenges selected by the bits of j.  out[j] (+)= sum_inst alpha_inst * h_inst[j]  (acc.rs:86-92 with m
// instances; m = 1, alpha = 1 for a single h).  A thread owns 16 consecutive coefficients: the product over the
// high bits is shared, the 16 low-bit combinations cost one multiplication each.
// xis: m x (lg_n+1) elements (xis[0] is xi_0, unused by h), alphas: m elements.
// ---------------------------------------------------------------------------------------------------------
template <int F>
__global__ void __launch_bounds__(128) ipa_hpoly_kernel(const Fe<F>* __restrict__ xis, const Fe<F>* __restrict__ alphas,
                                                        size_t m, int lg_n, Fe<F>* __restrict__ out) {
    const size_t n = (size_t)1 << lg_n;
    const size_t base = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 16;
    if (base >= n) return;
    const int lowbits = lg_n < 4 ? lg_n : 4;
    const int cnt = 1 << lowbits;
    Fe<F> acc[16];
#pragma unroll
    for (int t = 0; t < 16; ++t) acc[t] = fe_zero<F>();
    for (size_t inst = 0; inst < m; ++inst) {
        const Fe<F>* x = xis + inst * (size_t)(lg_n + 1);
        Fe<F> hi = g_ld(alphas + inst);
        for (int b = lowbits; b < lg_n; ++b)
            if ((base >> b) & 1) hi = fe_mul<F>(hi, g_ld(x + (lg_n - b)));
        // low 4 bits: v[t] for t = 0..15 built by doubling, v[t | 2^b] = v[t] * xi_{lg n - b}
        Fe<F> v[16];
        v[0] = hi;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (b < lowbits) {
                Fe<F> xb = g_ld(x + (lg_n - b));
#pragma unroll
                for (int t = 0; t < (1 << b); ++t) v[t | (1 << b)] = fe_mul<F>(v[t], xb);
            }
        }
#pragma unroll
        for (int t = 0; t < 16; ++t) if (t < cnt) acc[t] = fe_add<F>(acc[t], v[t]);
    }
#pragma unroll
    for (int t = 0; t < 16; ++t) if (t < cnt) g_st(out + base + t, acc[t]);
}

// ---------------------------------------------------------------------------------------------------------
// host-side group helpers (O(lg n) work only)
// ---------------------------------------------------------------------------------------------------------
template <int C> int msm_device(const void* d_bases, const void* d_scalars, size_t k, size_t nbatch, uint64_t* out_xyz,
                                ptrdiff_t base_stride);

// fixed-base table for the many multiplications by H' inside one opening: T[i][d-1] = d * 16^i * P
template <int C>
struct FixedBase {
    typedef Host<C> H;
    std::vector<typename H::Pt> tab;
    explicit FixedBase(const typename H::Pt& P) : tab(64 * 15) {
        typename H::Pt base = P;
        for (int i = 0; i < 64; ++i) {
            tab[i * 15] = base;
            for (int d = 2; d <= 15; ++d) tab[i * 15 + d - 1] = H::add(tab[i * 15 + d - 2], base);
            base = H::add(tab[i * 15 + 14], base);       // 16 * base
        }
    }
    typename H::Pt mul(const typename H::SF& k) const {
        Bits256 b = H::bits(k);
        typename H::Pt acc = xyzz_identity<C>();
        for (int i = 0; i < 64; ++i) {
            u32 d = (b.w[i >> 3] >> ((i & 7) * 4)) & 15u;
            if (d) xyzz_add<C>(acc, tab[i * 15 + d - 1]);
        }
        return acc;
    }
};

namespace glv_data {
#include "glv_consts.inc"
}
typedef unsigned __int128 u128;
// 4x4-limb schoolbook product
static void mul_4x4(const u64 a[4], const u64 b[4], u64 out[8]) {
    for (int i = 0; i < 8; ++i) out[i] = 0;
    for (int i = 0; i < 4; ++i) {
        u128 c = 0;
        for (int j = 0; j < 4; ++j) { c += (u128)a[i] * b[j] + out[i + j]; out[i + j] = (u64)c; c >>= 64; }
        out[i + 4] = (u64)c;
    }
}
// (a*b) mod 2^256, a - b mod 2^256
static void mul_lo4(const u64 a[4], const u64 b[4], u64 out[4]) { u64 t[8]; mul_4x4(a, b, t); for (int i = 0; i < 4; ++i) out[i] = t[i]; }
static void sub4(u64 a[4], const u64 b[4]) { u64 br = 0; for (int i = 0; i < 4; ++i) { u128 d = (u128)a[i] - b[i] - br; a[i] = (u64)d; br = (u64)(d >> 64) & 1; } }
static void neg4(u64 a[4]) { u64 z[4] = {0, 0, 0, 0}; sub4(z, a); for (int i = 0; i < 4; ++i) a[i] = z[i]; }

// xi = k1 + k2*lambda (mod r), |k1|, |k2| < 2^130.  Returns false (caller falls back to the plain ladder) if the
// split does not verify -- it always does for a correct constant table, the check is a guard, not a code path.
template <int C>
static bool glv_decompose(const Fe<CurveTraits<C>::SCALAR>& xi, GlvScalar* out) {
    constexpr int S = CurveTraits<C>::SCALAR;
    Fe<S> can = fe_to_canonical<S>(xi);
    u64 k[4]; memcpy(k, can.l, 32);
    u64 c[2][4];
    const unsigned long long* G[2] = {glv_data::GLV_G1[C], glv_data::GLV_G2[C]};
    for (int t = 0; t < 2; ++t) {                              // c_t = (k * g_t) >> 376
        u64 prod[8]; mul_4x4(k, (const u64*)G[t], prod);
        c[t][0] = (prod[5] >> 56) | (prod[6] << 8);
        c[t][1] = (prod[6] >> 56) | (prod[7] << 8);
        c[t][2] = prod[7] >> 56;
        c[t][3] = 0;
    }
    u64 k1[4], k2[4], t1[4], t2[4];
    memcpy(k1, k, 32);
    mul_lo4(c[0], (const u64*)glv_data::GLV_A1[C], t1); sub4(k1, t1);
    mul_lo4(c[1], (const u64*)glv_data::GLV_A2[C], t2); sub4(k1, t2);
    for (int i = 0; i < 4; ++i) k2[i] = 0;
    mul_lo4(c[0], (const u64*)glv_data::GLV_B1[C], t1); sub4(k2, t1);
    mul_lo4(c[1], (const u64*)glv_data::GLV_B2[C], t2); sub4(k2, t2);
    out->neg1 = (int)(k1[3] >> 63); if (out->neg1) neg4(k1);
    out->neg2 = (int)(k2[3] >> 63); if (out->neg2) neg4(k2);
    if (k1[3] || k2[3] || (k1[2] >> 32) || (k2[2] >> 32)) return false;
    // verify k1 + k2*lambda == xi in the scalar field
    auto to_fe = [](const u64 mag[4], int neg) {
        Fe<S> v; memcpy(v.l, mag, 32);
        v = fe_from_canonical<S>(v);
        return neg ? fe_neg<S>(v) : v;
    };
    Fe<S> lam; memcpy(lam.l, glv_data::GLV_LAMBDA[C], 32);
    lam = fe_from_canonical<S>(lam);
    if (!fe_eq<S>(fe_add<S>(to_fe(k1, out->neg1), fe_mul<S>(to_fe(k2, out->neg2), lam)), xi)) return false;
    memcpy(out->k1, k1, 20); memcpy(out->k2, k2, 20);
    // non-adjacent forms: k = sum d_i 2^i, d_i in {-1, 0, 1}, no two adjacent non-zero digits
    auto naf = [](const u64 mag[4], u32 pos[5], u32 neg[5]) {
        u128 k = ((u128)mag[1] << 64) | mag[0];
        u64 top = mag[2];                                  // bits 128.. (< 2^32)
        for (int i = 0; i < 5; ++i) pos[i] = neg[i] = 0;
        int i = 0;
        while ((k != 0 || top != 0) && i < 160) {
            if (k & 1) {
                if ((k & 3) == 1) { pos[i >> 5] |= 1u << (i & 31); k -= 1; }
                else { neg[i >> 5] |= 1u << (i & 31); k += 1; if (k == 0) top += 1; }     // carry out of the low 128 bits
            }
            k = (k >> 1) | ((u128)(top & 1) << 127);
            top >>= 1;
            ++i;
        }
        return i;
    };
    int l1 = naf(k1, out->p1, out->n1), l2 = naf(k2, out->p2, out->n2);
    out->nbits = l1 > l2 ? l1 : l2;
    return true;
}

static inline unsigned ilog2(size_t n) { unsigned l = 0; while (((size_t)1 << (l + 1)) <= n) ++l; return l; }
static inline size_t proof_words(unsigned lg_n) { return 24 * (size_t)lg_n + 33; }
static inline size_t instance_words(unsigned lg_n) { return 24 * (size_t)lg_n + 54; }

// trailing zero coefficients do not count (ark-poly's DensePolynomial strips them)
static size_t strip_len(const uint64_t* coeffs, size_t len) {
    while (len > 0 && !(coeffs[4 * len - 1] | coeffs[4 * len - 2] | coeffs[4 * len - 3] | coeffs[4 * len - 4])) --len;
    return len;
}

// sum of `count` per-block partials (device) -> host element
template <int F>
static int sum_partials(const Fe<F>* d_part, size_t count, Fe<F>* out, int nout) {
    Runtime& R = rt();
    void* h;
    if (int e = pinned_get(count * nout * 32, &h)) return e;
    Fe<F>* hp = (Fe<F>*)h;
    HALO_CUDA(cudaMemcpyAsync(hp, d_part, count * nout * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    for (int o = 0; o < nout; ++o) {
        Fe<F> s = fe_zero<F>();
        for (size_t i = 0; i < count; ++i) s = fe_add<F>(s, hp[i * nout + o]);
        out[o] = s;
    }
    return HALO_OK;
}

// ---------------------------------------------------------------------------------------------------------
// pcdl::commit (pcdl.rs:275-287) = pedersen::commit(w, Gs[0..n], coeffs) (pedersen.rs:7-27)
// ---------------------------------------------------------------------------------------------------------
template <int C>
static int commit_dev(const void* d_coeffs, size_t len, const typename Host<C>::SF* w, typename Host<C>::Pt* out) {
    typedef Host<C> H;
    Runtime& R = rt();
    uint64_t xyz[12];
    if (int e = msm_device<C>(R.pp[C].d_gs, d_coeffs, len, 1, xyz, 0)) return e;
    *out = H::from_words(xyz);
    if (w) *out = H::add(H::mul(H::from_affine_words(R.pp[C].s_xy), *w), *out);
    return HALO_OK;
}

template <int C>
int pcdl_commit(const uint64_t* coeffs, size_t len, size_t d, const uint64_t* w, uint64_t* out_xyz) {
    typedef Host<C> H;
    constexpr int S = H::S;
    Runtime& R = rt();
    const size_t n = d + 1;
    len = strip_len(coeffs, len);
    HALO_CHECK((n & (n - 1)) == 0, HALO_EINVAL, "commit: n is not a power of two");
    HALO_CHECK(len <= n, HALO_EINVAL, "commit: p_deg <= d violated");
    HALO_CHECK(R.pp[C].n > 0, HALO_EPARAMS, "commit: public parameters not loaded");
    HALO_CHECK(d <= R.pp[C].n - 1, HALO_EINVAL, "commit: d <= D violated");
    void* d_c = nullptr;
    if (len) {
        if (int e = ws_get(WS_SCALARS, len * 32, &d_c)) return e;
        HALO_CUDA(cudaMemcpyAsync(d_c, coeffs, len * 32, cudaMemcpyHostToDevice, R.stream));
    }
    typename H::SF ws;
    if (w) ws = H::scalar(w);
    typename H::Pt out;
    if (int e = commit_dev<C>(d_c, len, w ? &ws : nullptr, &out)) return e;
    H::to_words(out, out_xyz);
    (void)S;
    return HALO_OK;
}

// pcdl::chunked_commit (pcdl.rs:294-314): one pedersen::commit(w, Gs[0..n], chunk) per chunk of the (stripped)
// coefficient vector.  The full-length chunks are consecutive scalar vectors over the same bases, i.e. exactly the
// layout of the batched MSM; a ragged last chunk is one more MSM.  *out_count = number of chunks = ceil(len / chunk).
template <int C>
int pcdl_chunked_commit(const uint64_t* coeffs, size_t len, size_t d, const uint64_t* w, size_t chunk, uint64_t* out_xyz,
                        size_t* out_count) {
    typedef Host<C> H;
    Runtime& R = rt();
    const size_t n = d + 1;
    len = strip_len(coeffs, len);
    HALO_CHECK(n != 0 && (n & (n - 1)) == 0, HALO_EINVAL, "chunked_commit: n is not a power of two");
    HALO_CHECK(R.pp[C].n > 0, HALO_EPARAMS, "chunked_commit: public parameters not loaded");
    HALO_CHECK(d <= R.pp[C].n - 1, HALO_EINVAL, "chunked_commit: d <= D violated");
    HALO_CHECK(chunk > 0, HALO_EINVAL, "chunked_commit: chunk size must be non-zero (slice::chunks panics)");
    HALO_CHECK(len == 0 || std::min(chunk, len) <= n, HALO_EINVAL, "chunked_commit: a chunk is longer than Gs[0..n] (pedersen.rs:14-19)");
    const size_t full = len / chunk, rem = len - full * chunk, count = full + (rem ? 1 : 0);
    if (out_count) *out_count = count;
    if (!count) return HALO_OK;
    void* d_c;
    if (int e = ws_get(WS_SCALARS, len * 32, &d_c)) return e;
    HALO_CUDA(cudaMemcpyAsync(d_c, coeffs, len * 32, cudaMemcpyHostToDevice, R.stream));
    // batches bounded by the pipeline's 2^32 (window, point) pairs and a 1 GiB index buffer
    const size_t per = std::max<size_t>(1, std::min<size_t>(64, ((size_t)1 << 28) / (chunk * 17)));
    for (size_t i = 0; i < full; i += per) {
        const size_t b = std::min(per, full - i);
        if (int e = msm_device<C>(R.pp[C].d_gs, (const char*)d_c + i * chunk * 32, chunk, b, out_xyz + 12 * i, 0)) return e;
    }
    if (rem) if (int e = msm_device<C>(R.pp[C].d_gs, (const char*)d_c + full * chunk * 32, rem, 1, out_xyz + 12 * full, 0)) return e;
    if (w) {
        const typename H::Pt wS = H::mul(H::from_affine_words(R.pp[C].s_xy), H::scalar(w));
        for (size_t i = 0; i < count; ++i) H::to_words(H::add(wS, H::from_words(out_xyz + 12 * i)), out_xyz + 12 * i);
    }
    return HALO_OK;
}
template int pcdl_chunked_commit<0>(const uint64_t*, size_t, size_t, const uint64_t*, size_t, uint64_t*, size_t*);
template int pcdl_chunked_commit<1>(const uint64_t*, size_t, size_t, const uint64_t*, size_t, uint64_t*, size_t*);

// width-3 NAF of the two half-scalars as a list of table additions; false if a digit falls outside the table
static void wnaf3_ops(const u32 mag[5], int negative, u32 phi_bit, FoldOps* ops, bool* ok) {
    u64 k[3] = {(u64)mag[0] | ((u64)mag[1] << 32), (u64)mag[2] | ((u64)mag[3] << 32), (u64)mag[4]};
    for (int pos = 0; k[0] | k[1] | k[2]; ++pos) {
        if (k[0] & 1) {
            int d = (int)(k[0] & 7);                        // odd: 1, 3, 5, 7  ->  1, 3, -3, -1
            if (d > 4) d -= 8;
            if (d > 0) { u64 b0 = k[0]; k[0] -= (u64)d; if (k[0] > b0) { if (k[1]-- == 0) --k[2]; } }
            else { u64 b0 = k[0]; k[0] += (u64)(-d); if (k[0] < b0) { if (++k[1] == 0) ++k[2]; } }
            if (pos >= FOLD_ROWS || ops->n >= 160) { *ok = false; return; }
            const u32 row = (u32)pos + ((d == 3 || d == -3) ? (u32)FOLD_ROWS : 0u);
            const bool neg = (d < 0) != (negative != 0);
            ops->op[ops->n++] = (unsigned short)(row | phi_bit | (neg ? 0x400u : 0u));
        }
        k[0] = (k[0] >> 1) | (k[1] << 63); k[1] = (k[1] >> 1) | (k[2] << 63); k[2] >>= 1;
    }
}
static bool fold_ops_from(const GlvScalar& g, FoldOps* ops) {
    bool ok = true;
    ops->n = 0;
    wnaf3_ops(g.k1, g.neg1, 0u, ops, &ok);
    if (ok) wnaf3_ops(g.k2, g.neg2, 0x200u, ops, &ok);
    return ok;
}

// The fold table of Gs[n/2 .. n) for openings of size n, built on first use and kept with the parameter set
// (halo_pp_reset / halo_shutdown free it).  Rows are produced FOLD_SLICE at a time so that the XYZZ staging buffer
// stays at a quarter of the table's size; the affine conversion is the batched one the folds use.
template <int C>
static int build_fold_table(const Affine<C>* hi, size_t m, void** tab_out) {
    // rows 2^k * hi[j] and 3 * 2^k * hi[j], k < FOLD_ROWS, j < m; *tab_out == nullptr when the device has no room (ladder fold)
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    *tab_out = nullptr;
    void* tab = nullptr;
    if (cudaMalloc(&tab, (size_t)2 * FOLD_ROWS * m * sizeof(Affine<C>)) != cudaSuccess) { cudaGetLastError(); return HALO_OK; }
    XYZZ<C>*tmp, *state;
    if (int e = ws_get(WS_FOLD_TMP, (size_t)2 * FOLD_SLICE * m * sizeof(XYZZ<C>), (void**)&tmp)) { cudaFree(tab); return e; }
    if (int e = ws_get(WS_FOLD_STATE, m * sizeof(XYZZ<C>), (void**)&state)) { cudaFree(tab); return e; }
    for (int k0 = 0; k0 < FOLD_ROWS; k0 += FOLD_SLICE) {
        const size_t cnt = (size_t)FOLD_SLICE * m;
        ipa_fold_table_rows_kernel<C><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(hi, state, m, k0 == 0, FOLD_SLICE, tmp, tmp + cnt);
        for (int t3 = 0; t3 < 2; ++t3) {                   // rows of 2^k * G, then rows of 3 * 2^k * G
            Affine<C>* dst = (Affine<C>*)tab + ((size_t)t3 * FOLD_ROWS + k0) * m;
            const XYZZ<C>* src = tmp + (size_t)t3 * cnt;
            if (cnt >= 8192) xyzz_batch_affine_kernel<C, 8><<<(unsigned)((cnt / 8 + 128) / 128), 128, 0, st>>>(src, cnt, dst);
            else xyzz_batch_affine_kernel<C, 1><<<(unsigned)((cnt + 127) / 128), 128, 0, st>>>(src, cnt, dst);
        }
        R.launches += 3;
    }
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
        cudaFree(tab);
        last_err() = "fold table: kernel failed";
        return HALO_ECUDA;
    }
    *tab_out = tab;
    return HALO_OK;
}
template <int C>
static int ensure_fold_table(size_t n, const Affine<C>** out) {
    Runtime& R = rt();
    ResidentParams& P = R.pp[C];
    *out = nullptr;
    if (P.d_fold_tab && P.fold_tab_n == n) { *out = (const Affine<C>*)P.d_fold_tab; return HALO_OK; }
    if (R.is_lane) return HALO_OK;                         // a lane only borrows the main runtime's table: ladder fold without one
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    if (P.d_fold_tab) { cudaFree(P.d_fold_tab); P.d_fold_tab = nullptr; P.fold_tab_n = 0; }
    void* tab = nullptr;
    if (int e = build_fold_table<C>((const Affine<C>*)P.d_gs + n / 2, n / 2, &tab)) return e;
    if (!tab) return HALO_OK;
    P.d_fold_tab = tab; P.fold_tab_n = n;
    *out = (const Affine<C>*)tab;
    return HALO_OK;
}
// the same table over THIS RANK's generators of a sharded opening of size n (gl_hi: the upper half of the rank's
// interleaved subset of pp.Gs, n / (2 world) points), kept with the parameter set like the global one
template <int C>
static int ensure_fold_table_local(size_t n, int world, int rank, const Affine<C>* gl_hi, const Affine<C>** out) {
    Runtime& R = rt();
    ResidentParams& P = R.pp[C];
    const int s = dist_in_half() ? 1 : 0;
    *out = nullptr;
    if (P.d_fold_tab_loc[s] && P.fold_tab_loc_n[s] == n && P.fold_tab_loc_world[s] == world && P.fold_tab_loc_rank[s] == rank) {
        *out = (const Affine<C>*)P.d_fold_tab_loc[s]; return HALO_OK;
    }
    if (R.is_lane) return HALO_OK;
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    if (P.d_fold_tab_loc[s]) { cudaFree(P.d_fold_tab_loc[s]); P.d_fold_tab_loc[s] = nullptr; P.fold_tab_loc_n[s] = 0; }
    void* tab = nullptr;
    if (int e = build_fold_table<C>(gl_hi, n / 2 / (size_t)world, &tab)) return e;
    if (!tab) return HALO_OK;
    P.d_fold_tab_loc[s] = tab; P.fold_tab_loc_n[s] = n; P.fold_tab_loc_world[s] = world; P.fold_tab_loc_rank[s] = rank;
    *out = (const Affine<C>*)tab;
    return HALO_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Fixed-base table of the resident generators for the COMMIT path (msm.cu, msm_fixed): row w holds 2^(c w) * Gs[i],
// i < pp.n, affine.  With it the digit columns of a commitment all land in ONE window of 2^(c-1) buckets (the bucket
// reduction and the Horner combine all but disappear) and c can be 18-20 instead of 15-16: 13 additions per point
// instead of 16 at n = 2^20.  Built once per parameter set on first use (13 x 64 MB at 2^20, ~40 ms), freed with it.
// ---------------------------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(128) fixed_table_step_kernel(const Affine<C>* __restrict__ gs, XYZZ<C>* __restrict__ state, size_t n,
                                                               int first, int doublings) {
    size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (j >= n) return;
    XYZZ<C> P;
    if (first) { Affine<C> a; a.x = g_ld(&gs[j].x); a.y = g_ld(&gs[j].y); P = xyzz_from_affine<C>(a); }
    else { P.X = g_ld(&state[j].X); P.Y = g_ld(&state[j].Y); P.ZZ = g_ld(&state[j].ZZ); P.ZZZ = g_ld(&state[j].ZZZ); }
    for (int k = 0; k < doublings; ++k) P = xyzz_dbl<C>(P);
    g_st(&state[j].X, P.X); g_st(&state[j].Y, P.Y); g_st(&state[j].ZZ, P.ZZ); g_st(&state[j].ZZZ, P.ZZZ);
}
template <int C>
int ensure_fixed_table(const void** tab_out, int* c_out, int* ncols_out, size_t* row_stride_out) {
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    ResidentParams& P = R.pp[C];
    *tab_out = nullptr;
    const size_t n = P.n;
    if (n < ((size_t)1 << 18) || n > ((size_t)1 << 24)) return HALO_OK;
    if (P.d_fixed_tab && P.fixed_tab_n == n) {
        *tab_out = P.d_fixed_tab; *c_out = P.fixed_tab_c; *ncols_out = 255 / P.fixed_tab_c + 1; *row_stride_out = n;
        return HALO_OK;
    }
    if (R.is_lane) return HALO_OK;
    HALO_CUDA(cudaStreamSynchronize(st));
    if (P.d_fixed_tab) { cudaFree(P.d_fixed_tab); P.d_fixed_tab = nullptr; P.fixed_tab_n = 0; }
    const int c = std::min(20, std::max(18, (int)ilog2(n)));
    const int ncols = 255 / c + 1;
    void* tab = nullptr;
    if (cudaMalloc(&tab, (size_t)ncols * n * sizeof(Affine<C>)) != cudaSuccess) { cudaGetLastError(); return HALO_OK; }   // no room: ordinary MSM
    XYZZ<C>* state;
    if (int e = ws_get(WS_FOLD_STATE, n * sizeof(XYZZ<C>), (void**)&state)) { cudaFree(tab); return e; }
    HALO_CUDA(cudaMemcpyAsync(tab, P.d_gs, n * sizeof(Affine<C>), cudaMemcpyDeviceToDevice, st));     // row 0 is Gs itself
    for (int w = 1; w < ncols; ++w) {
        fixed_table_step_kernel<C><<<(unsigned)((n + 127) / 128), 128, 0, st>>>((const Affine<C>*)P.d_gs, state, n, w == 1, c);
        xyzz_batch_affine_kernel<C, 8><<<(unsigned)((n / 8 + 128) / 128), 128, 0, st>>>(state, n, (Affine<C>*)tab + (size_t)w * n);
        R.launches += 2;
    }
    if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
        cudaFree(tab);
        last_err() = "fixed-base table: kernel failed";
        return HALO_ECUDA;
    }
    P.d_fixed_tab = tab; P.fixed_tab_n = n; P.fixed_tab_c = c;
    *tab_out = tab; *c_out = c; *ncols_out = ncols; *row_stride_out = n;
    return HALO_OK;
}
template int ensure_fixed_table<0>(const void**, int*, int*, size_t*);
template int ensure_fixed_table<1>(const void**, int*, int*, size_t*);

// explicit set-up step (halo_ipa_fold_table_build): build the table for openings of size n now instead of inside the
// first opening
template <int C>
int fold_table_build(size_t n) {
    Runtime& R = rt();
    HALO_CHECK(n >= 4 && (n & (n - 1)) == 0, HALO_EINVAL, "fold table: n must be a power of two >= 4");
    HALO_CHECK(R.pp[C].n >= n, HALO_EPARAMS, "fold table: public parameters missing or shorter than n");
    const Affine<C>* tab = nullptr;
    if (int e = ensure_fold_table<C>(n, &tab)) return e;
    HALO_CHECK(tab != nullptr, HALO_ECUDA, "fold table: not enough device memory (openings fall back to the ladder fold)");
    return HALO_OK;
}
template int fold_table_build<0>(size_t);
template int fold_table_build<1>(size_t);

// Interleaved ownership of a sharded opening: rank r of W holds the elements i = j * W + r.  E = uint4 words per element
// (2 for a scalar, 4 for an affine point).  take: out[j] = in[j * W + r];  weave: out[j * W + p] = in[p * cnt + j] (the
// all-gathered per-rank blocks back into index order).
__global__ void __launch_bounds__(256) ipa_take_kernel(const uint4* __restrict__ in, size_t cnt, unsigned W, unsigned r, unsigned E,
                                                       uint4* __restrict__ out) {
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= cnt * E) return;
    const size_t j = t / E; const unsigned w = (unsigned)(t - j * E);
    out[t] = in[(j * W + r) * E + w];
}
__global__ void __launch_bounds__(256) ipa_weave_kernel(const uint4* __restrict__ in, size_t cnt, unsigned W, unsigned E,
                                                        uint4* __restrict__ out) {
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= cnt * W * E) return;
    const size_t e = t / E; const unsigned w = (unsigned)(t - e * E);
    const size_t p = e / cnt, j = e - p * cnt;
    out[(j * W + p) * E + w] = in[t];
}

// fixed-base table for the first fold (tuning knob 0: 1 = never, 2 = for every size; by default for openings over the
// whole parameter set of at least 2^16 generators, the size a prover keeps using)
template <int C>
static bool open_uses_fold_table(size_t n) {
    Runtime& R = rt();
    const int tail_lg = R.tune[3] == 0 ? 15 : R.tune[3];
    const size_t tail_max = tail_lg < 0 ? 0 : (size_t)1 << tail_lg;
    return R.tune[0] != 1 && R.tune[2] != 1 && n >= 4 && n / 2 > tail_max && (R.tune[0] == 2 || (n == R.pp[C].n && n >= ((size_t)1 << 16)));
}
// what an opening of size n builds lazily on the MAIN runtime, done ahead of openings that run on a lane
template <int C>
int pcdl_open_prepare(size_t n) {
    const Affine<C>* tab = nullptr;
    if (open_uses_fold_table<C>(n)) return ensure_fold_table<C>(n, &tab);
    return HALO_OK;
}
template int pcdl_open_prepare<0>(size_t);
template int pcdl_open_prepare<1>(size_t);

// ---------------------------------------------------------------------------------------------------------
// pcdl::open / open_without_eval (pcdl.rs:326-473)
//   proof layout (u64 words): Ls[12 lg n] Rs[12 lg n] U[12] c[4] hiding[1] C_bar[12] w_prime[4]
// ---------------------------------------------------------------------------------------------------------
template <int C>
int pcdl_open(const uint64_t* coeffs, size_t len, const uint64_t* C_xyz, size_t d, const uint64_t* z_w,
              const uint64_t* v_w, const uint64_t* w_w, const uint64_t* hide_q, const uint64_t* hide_wbar,
              uint64_t* out_proof, uint64_t* out_v, bool coeffs_resident, bool shard) {
    // coeffs_resident: the n zero-padded coefficients are already in the WS_IPA_C workspace (acc::prover's h(X))
    // shard (collective; same inputs on every rank of the library's communicator): the rounds above the fold-free tail
    // run on interleaved slices -- rank r of W owns the indices i = j W + r of the generator, coefficient and power
    // vectors, so that a fold pairs (i, i + m) INSIDE a rank for every m that is a multiple of W.  Locally that is an
    // ordinary opening of size n / W (same kernels, same fold variants, a fold table over the rank's own generators),
    // except that the two dot products are summed over the ranks and L / R are point-range-sharded MSMs
    // (dist_msm_device).  When the tail starts (<= 2^15 generators, latency-bound) the three vectors are all-gathered
    // and every rank finishes the opening redundantly.  Same transcript on every rank, same proof as the single-GPU path.
    typedef Host<C> H;
    typedef typename H::SF SF;
    typedef typename H::Pt Pt;
    constexpr int S = H::S;
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    const size_t n = d + 1;
    HALO_CHECK(n > 1 && (n & (n - 1)) == 0, HALO_EINVAL, "open: n is not a power of two (or n <= 1)");
    if (!coeffs_resident) len = strip_len(coeffs, len);
    HALO_CHECK(len <= n, HALO_EINVAL, "open: p.degree() <= d violated");
    HALO_CHECK(R.pp[C].n > 0, HALO_EPARAMS, "open: public parameters not loaded");
    HALO_CHECK(d <= R.pp[C].n - 1, HALO_EINVAL, "open: d <= pp.D violated");
    HALO_CHECK(!w_w || (hide_q && hide_wbar), HALO_EINVAL, "open: hiding needs the prover's randomness (q, w_bar)");
    const unsigned lg_n = ilog2(n);
    const SF z = H::scalar(z_w);

    Affine<C>* gs; SF *cs, *zs, *part, *tmp;
    if (int e = ws_get(WS_IPA_G, n * 64, (void**)&gs)) return e;
    if (int e = ws_get(WS_IPA_C, n * 32, (void**)&cs)) return e;
    if (int e = ws_get(WS_IPA_Z, n * 32, (void**)&zs)) return e;
    if (int e = ws_get(WS_IPA_T0, 2 * 256 * 32 + 64, (void**)&part)) return e;
    // cs = coefficients of p, zero-padded to n (pcdl.rs:392-393)
    if (!coeffs_resident) {
        if (len) HALO_CUDA(cudaMemcpyAsync(cs, coeffs, len * 32, cudaMemcpyHostToDevice, st));
        if (len < n) HALO_CUDA(cudaMemsetAsync(cs + len, 0, (n - len) * 32, st));
    }
    // zs = (1, z, ..., z^(n-1)) (pcdl.rs:395)
    ipa_powers_kernel<S><<<(unsigned)((n / 16 + 128) / 128), 128, 0, st>>>(zs, n, z, fe_one<S>());
    R.launches += 1;
    HALO_CUDA(cudaGetLastError());

    // v = p(z) = <cs, zs> when the caller did not supply it (pcdl::open, pcdl.rs:471)
    SF v;
    if (v_w) v = H::scalar(v_w);
    else {
        const int g = (int)std::min<size_t>(256, (n + 255) / 256);
        ipa_dot1_kernel<S><<<g, 256, 0, st>>>(cs, zs, n, part);
        R.launches += 1;
        if (int e = sum_partials<S>(part, g, &v, 1)) return e;
    }
    if (out_v) H::put_scalar(v, out_v);

    Transcript<C> tr(LABEL_PCDL);
    Pt Cp = H::from_words(C_xyz);
    const unsigned L = lg_n;
    uint64_t* o_hiding = out_proof + 24 * L + 16;
    memset(o_hiding, 0, 17 * 8);
    if (w_w) {
        // p_bar = q (X - z), C_bar = commit(p_bar; w_bar), a = rho(C, C_bar, z, v), p' = p + a p_bar,
        // w' = w + a w_bar, C' = C + a C_bar - w' S   (pcdl.rs:344-371)
        const SF w = H::scalar(w_w), wbar = H::scalar(hide_wbar);
        SF* q;
        if (int e = ws_get(WS_IPA_T1, 2 * n * 32, (void**)&q)) return e;
        tmp = q + n;
        HALO_CUDA(cudaMemcpyAsync(q, hide_q, d * 32, cudaMemcpyHostToDevice, st));
        ipa_pbar_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(q, n, z, tmp);
        R.launches += 1;
        Pt Cbar;
        if (int e = commit_dev<C>(tmp, n, &wbar, &Cbar)) return e;
        tr.absorb_point(Cp); tr.absorb_point(Cbar);
        tr.absorb_scalar(z); tr.absorb_scalar(v);
        const SF a = tr.challenge();
        ipa_axpy_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(cs, tmp, n, a);
        R.launches += 1;
        const SF wprime = fe_add<S>(fe_mul<S>(wbar, a), w);
        Cp = H::add(H::add(Cp, H::mul(Cbar, a)), H::neg(H::mul(H::from_affine_words(R.pp[C].s_xy), wprime)));
        o_hiding[0] = 1;
        H::to_words(Cbar, o_hiding + 1);
        H::put_scalar(wprime, o_hiding + 13);
    }
    tr.absorb_point(Cp);
    tr.absorb_scalar(z); tr.absorb_scalar(v);
    SF xi = tr.challenge();
    FixedBase<C> Hp(H::mul(H::from_affine_words(R.pp[C].h_xy), xi));

    // sharded rounds need a power-of-two world that leaves every rank a whole tail's worth of work above the tail
    int W = 1, rank = 0;
    {
        const int tl = R.tune[3] == 0 ? 15 : R.tune[3];
        if (shard && dist_on() && dist_world() > 1 && (dist_world() & (dist_world() - 1)) == 0 && tl >= 8 && tl <= 20 &&
            n >= ((size_t)4 << tl) && (size_t)dist_world() <= ((size_t)1 << (tl - 2)) && !R.is_lane) {
            W = dist_world(); rank = dist_rank();
        }
    }
    Affine<C>* gs_full = gs; SF *cs_full = cs, *zs_full = zs;       // the replicated tail continues in these
    if (W == 1) {
        // gs = pp.Gs[0..n] (pcdl.rs:394)
        HALO_CUDA(cudaMemcpyAsync(gs, R.pp[C].d_gs, n * 64, cudaMemcpyDeviceToDevice, st));
    } else {
        const size_t nl = n / (size_t)W;
        Affine<C>* gl; SF *cl, *zl;
        if (int e = ws_get(WS_IPA_GL, nl * 64, (void**)&gl)) return e;
        if (int e = ws_get(WS_IPA_CL, 2 * nl * 32, (void**)&cl)) return e;
        zl = cl + nl;
        ipa_take_kernel<<<(unsigned)((nl * 4 + 255) / 256), 256, 0, st>>>((const uint4*)R.pp[C].d_gs, nl, (unsigned)W, (unsigned)rank, 4, (uint4*)gl);
        ipa_take_kernel<<<(unsigned)((nl * 2 + 255) / 256), 256, 0, st>>>((const uint4*)cs, nl, (unsigned)W, (unsigned)rank, 2, (uint4*)cl);
        ipa_take_kernel<<<(unsigned)((nl * 2 + 255) / 256), 256, 0, st>>>((const uint4*)zs, nl, (unsigned)W, (unsigned)rank, 2, (uint4*)zl);
        R.launches += 3;
        HALO_CUDA(cudaGetLastError());
        gs = gl; cs = cl; zs = zl;
    }

    // optional phase trace (tuning knob 7): wall clock with a stream sync at each phase boundary
    const bool trace = R.tune[7] != 0;
    double t_dots = 0, t_msm = 0, t_host = 0, t_fold = 0;
    auto now = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    // fold-free tail below 2^tail_lg generators (tuning knob 3: 0 = default 15, -1 = never)
    const int tail_lg = R.tune[3] == 0 ? 15 : R.tune[3];
    const size_t tail_max = tail_lg < 0 ? 0 : (size_t)1 << tail_lg;
    // radix-4 generator folds above the tail (tuning knob 4 = 1 disables; the literal-ladder knob disables it too)
    const bool radix4 = R.tune[4] != 1 && R.tune[2] != 1;
    // fixed-base table for the first fold (tuning knob 0: 1 = never, 2 = for every size; by default for openings
    // over the whole parameter set of at least 2^16 generators, the size a prover keeps using)
    const Affine<C>* fold_tab = nullptr;
    if (open_uses_fold_table<C>(n)) {
        if (W == 1) { if (int e = ensure_fold_table<C>(n, &fold_tab)) return e; }
        else if (int e = ensure_fold_table_local<C>(n, W, rank, gs + n / 2 / (size_t)W, &fold_tab)) return e;
    }
    bool pending = false;
    SF pend_xi = fe_zero<S>();
    bool tail = false;
    size_t T = 0, nu = 0;
    int w_cur = 0;
    SF *tail_s = nullptr, *tail_w = nullptr;
    size_t m = n / 2 / (size_t)W;                         // LOCAL half-length while the rounds are sharded (W > 1)
    for (unsigned r = 0; r < lg_n; ++r, m /= 2) {
        double t0 = trace ? now() : 0;
        if (W > 1 && !pending && 2 * m * (size_t)W <= tail_max) {
            // the tail starts: gather the interleaved slices (2m elements per rank) back into index order; from here on
            // every rank holds the whole (small) problem and the rounds run as on one GPU
            const size_t cnt = 2 * m;
            unsigned char* stage;
            if (int e = ws_get(WS_IPA_T1, (size_t)W * cnt * 128, (void**)&stage)) return e;
            unsigned char* sg = stage; unsigned char* sc = stage + (size_t)W * cnt * 64; unsigned char* sz = sc + (size_t)W * cnt * 32;
            HALO_CUDA(cudaMemcpyAsync(sg + (size_t)rank * cnt * 64, gs, cnt * 64, cudaMemcpyDeviceToDevice, st));
            HALO_CUDA(cudaMemcpyAsync(sc + (size_t)rank * cnt * 32, cs, cnt * 32, cudaMemcpyDeviceToDevice, st));
            HALO_CUDA(cudaMemcpyAsync(sz + (size_t)rank * cnt * 32, zs, cnt * 32, cudaMemcpyDeviceToDevice, st));
            if (int e = dist_allgather_dev(sg + (size_t)rank * cnt * 64, sg, cnt * 64)) return e;
            if (int e = dist_allgather_dev(sc + (size_t)rank * cnt * 32, sc, cnt * 32)) return e;
            if (int e = dist_allgather_dev(sz + (size_t)rank * cnt * 32, sz, cnt * 32)) return e;
            ipa_weave_kernel<<<(unsigned)((cnt * W * 4 + 255) / 256), 256, 0, st>>>((const uint4*)sg, cnt, (unsigned)W, 4, (uint4*)gs_full);
            ipa_weave_kernel<<<(unsigned)((cnt * W * 2 + 255) / 256), 256, 0, st>>>((const uint4*)sc, cnt, (unsigned)W, 2, (uint4*)cs_full);
            ipa_weave_kernel<<<(unsigned)((cnt * W * 2 + 255) / 256), 256, 0, st>>>((const uint4*)sz, cnt, (unsigned)W, 2, (uint4*)zs_full);
            R.launches += 3;
            HALO_CUDA(cudaGetLastError());
            gs = gs_full; cs = cs_full; zs = zs_full;
            m *= (size_t)W; W = 1;
        }
        const int g = (int)std::min<size_t>(256, (m + 255) / 256);
        ipa_dots_kernel<S><<<g, 256, 0, st>>>(cs, zs, m, part);
        R.launches += 1;
        SF dots[2];
        if (int e = sum_partials<S>(part, g, dots, 2)) return e;
        if (W > 1) {                                       // the dot products are sums over all ranks' slices
            std::vector<SF> all((size_t)2 * W);
            if (int e = dist_allgather_host(dots, all.data(), 64)) return e;
            dots[0] = fe_zero<S>(); dots[1] = fe_zero<S>();
            for (int p = 0; p < W; ++p) { dots[0] = fe_add<S>(dots[0], all[2 * p]); dots[1] = fe_add<S>(dots[1], all[2 * p + 1]); }
        }
        // one pipeline, two MSMs: [0] = R = <l(c), r(G)>, [1] = L = <r(c), l(G)>  (pcdl.rs:412-418)
        uint64_t lr[24];
        double t1 = trace ? now() : 0;
        if (!tail && !pending && 2 * m <= tail_max && W == 1) {     // switch: freeze the generators, start the weights at [1]
            tail = true; T = 2 * m; nu = 1;
            if (int e = ws_get(WS_IPA_S, 2 * T * 32, (void**)&tail_s)) return e;
            if (int e = ws_get(WS_IPA_W, 2 * T * 32, (void**)&tail_w)) return e;
            const SF one = fe_one<S>();
            HALO_CUDA(cudaMemcpyAsync(tail_w, &one, 32, cudaMemcpyHostToDevice, st));
            w_cur = 0;
        }
        if (tail) {
            ipa_tail_scalars_kernel<S><<<(unsigned)((T + 255) / 256), 256, 0, st>>>(cs, tail_w + w_cur * T, m, T, tail_s, tail_s + T);
            R.launches += 1;
            if (int e = msm_device<C>(gs, tail_s, T, 2, lr, 0)) return e;
        } else if (pending) {                             // one fold is pending: generators of length 4m with weights [1, xi_a]
            const size_t Tp = 4 * m;
            if (int e = ws_get(WS_IPA_S, 2 * Tp * 32, (void**)&tail_s)) return e;
            if (int e = ws_get(WS_IPA_W, 2 * 32, (void**)&tail_w)) return e;
            const SF wh[2] = {fe_one<S>(), pend_xi};
            HALO_CUDA(cudaMemcpyAsync(tail_w, wh, 64, cudaMemcpyHostToDevice, st));
            ipa_tail_scalars_kernel<S><<<(unsigned)((Tp + 255) / 256), 256, 0, st>>>(cs, tail_w, m, Tp, tail_s, tail_s + Tp);
            R.launches += 1;
            if (W > 1) { if (int e = dist_msm_device<C>(gs, tail_s, Tp, 2, lr, 0)) return e; }
            else if (int e = msm_device<C>(gs, tail_s, Tp, 2, lr, 0)) return e;     // syncs: wh may go out of scope
        } else {
            if (W > 1) { if (int e = dist_msm_device<C>(gs + m, cs, m, 2, lr, -(ptrdiff_t)m)) return e; }
            else if (r == 0 && n == R.pp[C].n) {
                // the generators are still pp.Gs: addressed in place, msm_device takes the commit-path table for the pair
                if (int e = msm_device<C>((const Affine<C>*)R.pp[C].d_gs + m, cs, m, 2, lr, -(ptrdiff_t)m)) return e;
            }
            else if (int e = msm_device<C>(gs + m, cs, m, 2, lr, -(ptrdiff_t)m)) return e;
        }
        double t2 = trace ? now() : 0;
        Pt Rp = H::add(H::from_words(lr), Hp.mul(dots[1]));
        Pt Lp = H::add(H::from_words(lr + 12), Hp.mul(dots[0]));
        H::to_words(Lp, out_proof + 12 * r);
        H::to_words(Rp, out_proof + 12 * (L + r));
        // xi_{i+1} = rho(xi_i, L, R)  (pcdl.rs:421-425)
        tr.absorb_scalar(xi);
        tr.absorb_point(Lp); tr.absorb_point(Rp);
        const SF xi_next = tr.challenge();
        HALO_CHECK(!fe_is_zero<S>(xi_next), HALO_EINVAL, "open: zero challenge has no inverse");
        const SF xi_inv = fe_inv<S>(xi_next);
        xi = xi_next;
        double t3 = trace ? now() : 0;
        GlvScalar gk;
        FoldOps fops;
        if (tail) {
            const size_t th = std::max(m, nu);
            ipa_tail_fold_kernel<S><<<(unsigned)((th + 255) / 256), 256, 0, st>>>(cs, zs, m, tail_w + w_cur * T, nu,
                                                                                  tail_w + (1 - w_cur) * T, xi_next, xi_inv);
            R.launches += 1;
            w_cur = 1 - w_cur; nu *= 2;
        } else if (pending) {
            // second of two delayed folds: materialise g'[i] = B[i] + xi_b B[i+m] + xi_a B[i+2m] + xi_a xi_b B[i+3m]
            JointScalars js;
            const SF ws3[3] = {xi_next, pend_xi, fe_mul<S>(pend_xi, xi_next)};
            bool ok = true;
            js.nbits = 0;
            for (int t = 0; t < 3; ++t) { ok = ok && glv_decompose<C>(ws3[t], &js.s[t]); js.nbits = std::max(js.nbits, js.s[t].nbits); }
            HALO_CHECK(ok, HALO_EINVAL, "open: scalar split failed");
            ipa_tail_fold_kernel<S><<<(unsigned)((m + 255) / 256), 256, 0, st>>>(cs, zs, m, nullptr, 0, nullptr, xi_next, xi_inv);
            XYZZ<C>* folded;
            if (int e = ws_get(WS_IPA_T1, m * sizeof(XYZZ<C>), (void**)&folded)) return e;
            Fe<H::B> beta; memcpy(&beta, glv_data::GLV_BETA[C], 32);
            ipa_materialize_kernel<C, 3><<<(unsigned)((m + 127) / 128), 128, 9 * 2 * 128 * sizeof(uint4), st>>>(gs, m, js, beta, folded);
            if (m >= 8192) xyzz_batch_affine_kernel<C, 8><<<(unsigned)((m / 8 + 128) / 128), 128, 0, st>>>(folded, m, gs);
            else xyzz_batch_affine_kernel<C, 1><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(folded, m, gs);
            R.launches += 3;
            pending = false;
        } else if (fold_tab && r == 0 && glv_decompose<C>(xi_next, &gk) && fold_ops_from(gk, &fops)) {
            // the generators are still pp.Gs: xi * Gs[m + j] from the table, no doublings
            ipa_tail_fold_kernel<S><<<(unsigned)((m + 255) / 256), 256, 0, st>>>(cs, zs, m, nullptr, 0, nullptr, xi_next, xi_inv);
            XYZZ<C>* folded;
            if (int e = ws_get(WS_IPA_T1, m * sizeof(XYZZ<C>), (void**)&folded)) return e;
            Fe<H::B> beta; memcpy(&beta, glv_data::GLV_BETA[C], 32);
            ipa_fold_table_kernel<C><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(gs, fold_tab, m, fops, beta, folded);
            if (m >= 8192) xyzz_batch_affine_kernel<C, 8><<<(unsigned)((m / 8 + 128) / 128), 128, 0, st>>>(folded, m, gs);
            else xyzz_batch_affine_kernel<C, 1><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(folded, m, gs);
            R.launches += 3;
        } else if (radix4 && m * (size_t)W > tail_max && m >= 2) {
            // delay this fold: only the scalar vectors are folded now, the generators wait for the next challenge
            ipa_tail_fold_kernel<S><<<(unsigned)((m + 255) / 256), 256, 0, st>>>(cs, zs, m, nullptr, 0, nullptr, xi_next, xi_inv);
            R.launches += 1;
            pending = true; pend_xi = xi_next;
        } else if (R.tune[2] != 1 && glv_decompose<C>(xi_next, &gk)) {
            XYZZ<C>* folded;
            if (int e = ws_get(WS_IPA_T1, m * sizeof(XYZZ<C>), (void**)&folded)) return e;
            Fe<H::B> beta; memcpy(&beta, glv_data::GLV_BETA[C], 32);
            ipa_fold_glv_kernel<C><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(gs, cs, zs, m, gk, beta, xi_next, xi_inv, folded);
            if (m >= 8192) xyzz_batch_affine_kernel<C, 8><<<(unsigned)((m / 8 + 128) / 128), 128, 0, st>>>(folded, m, gs);
            else xyzz_batch_affine_kernel<C, 1><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(folded, m, gs);
            R.launches += 2;
        } else {                                        // literal ladder (tuning knob 2 = 1 forces it; parity tests use both)
            ipa_fold_kernel<C><<<(unsigned)((m + 127) / 128), 128, 0, st>>>(gs, cs, zs, m, H::bits(xi_next), xi_next, xi_inv);
            R.launches += 1;
        }
        HALO_CUDA(cudaGetLastError());
        if (trace) {
            HALO_CUDA(cudaStreamSynchronize(st));
            double t4 = now();
            t_dots += t1 - t0; t_msm += t2 - t1; t_host += t3 - t2; t_fold += t4 - t3;
            fprintf(stderr, "[halo trace] open round %u m=%zu: dots %.3f msm %.3f host %.3f fold %.3f ms\n", r, m, t1 - t0, t2 - t1, t3 - t2, t4 - t3);
        }
    }
    if (trace) fprintf(stderr, "[halo trace] open n=%zu totals: dots %.2f msm %.2f host %.2f fold %.2f ms\n", n, t_dots, t_msm, t_host, t_fold);
    // U = G_(lg n), c = c_(lg n)  (pcdl.rs:441-450)
    Affine<C> U; SF c;
    HALO_CUDA(cudaMemcpyAsync(&c, cs, 32, cudaMemcpyDeviceToHost, st));
    if (tail) {                                           // U = <w, B>
        uint64_t uw[12];
        if (int e = msm_device<C>(gs, tail_w + w_cur * T, T, 1, uw, 0)) return e;
        HALO_CUDA(cudaStreamSynchronize(st));
        memcpy(out_proof + 24 * L, uw, 96);
        H::put_scalar(c, out_proof + 24 * L + 12);
        return HALO_OK;
    }
    HALO_CUDA(cudaMemcpyAsync(&U, gs, 64, cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    H::to_words(xyzz_from_affine<C>(U), out_proof + 24 * L);
    H::put_scalar(c, out_proof + 24 * L + 12);
    return HALO_OK;
}

// ---------------------------------------------------------------------------------------------------------
// HPoly::eval (pcdl.rs:221-233): h(z) = prod_{i<lg n} (1 + xi_{lg n - i} z^{2^i})
// ---------------------------------------------------------------------------------------------------------
template <int S>
static Fe<S> hpoly_eval(const Fe<S>* xis, unsigned lg_n, const Fe<S>& z) {
    // literally as the reference: the first factor is taken before the loop, so lg n = 0 yields 1 + xi_0 z
    Fe<S> one = fe_one<S>(), zi = z;
    Fe<S> v = fe_add<S>(one, fe_mul<S>(xis[lg_n], z));
    for (unsigned i = 1; i < lg_n; ++i) {
        zi = fe_sqr<S>(zi);
        v = fe_mul<S>(v, fe_add<S>(one, fe_mul<S>(xis[lg_n - i], zi)));
    }
    return v;
}

// flat instance (u64 words): C[12] d[1] z[4] v[4] | proof (see pcdl_open)
template <int C>
struct InstanceView {
    typedef Host<C> H;
    const uint64_t* w;
    size_t d;
    unsigned lg_n;
    explicit InstanceView(const uint64_t* words) : w(words), d((size_t)words[12]), lg_n(ilog2((size_t)words[12] + 1)) {}
    typename H::Pt Cm() const { return H::from_words(w); }
    typename H::SF z() const { return H::scalar(w + 13); }
    typename H::SF v() const { return H::scalar(w + 17); }
    const uint64_t* proof() const { return w + 21; }
    typename H::Pt Lr(unsigned i) const { return H::from_words(proof() + 12 * i); }
    typename H::Pt Rr(unsigned i) const { return H::from_words(proof() + 12 * (lg_n + i)); }
    typename H::Pt U() const { return H::from_words(proof() + 24 * lg_n); }
    typename H::SF c() const { return H::scalar(proof() + 24 * lg_n + 12); }
    bool hiding() const { return proof()[24 * lg_n + 16] != 0; }
    typename H::Pt Cbar() const { return H::from_words(proof() + 24 * lg_n + 17); }
    typename H::SF wprime() const { return H::scalar(proof() + 24 * lg_n + 29); }
};

// What the reference gets from its types -- arkworks field elements are always below the modulus and deserialised
// points are checked to be on the curve -- has to be checked at the C ABI: a flat instance is raw words.  `d` is
// validated FIRST, because every other offset is derived from it.  HALO_EINVAL for a malformed d (n not a power of
// two, as the assert in the reference), HALO_EREJECT for d > D, a non-canonical scalar or an off-curve point.
template <int C>
static int instance_validate(const uint64_t* inst) {
    typedef Host<C> H;
    Runtime& R = rt();
    const uint64_t d = inst[12];
    HALO_CHECK(d != ~(uint64_t)0 && ((d + 1) & d) == 0, HALO_EINVAL, "instance: n = d + 1 is not a power of two");
    HALO_CHECK(R.pp[C].n > 0, HALO_EPARAMS, "instance: public parameters not loaded");
    HALO_CHECK(d <= R.pp[C].n - 1, HALO_EREJECT, "d was larger than D!");
    InstanceView<C> q(inst);
    const unsigned L = q.lg_n;
    HALO_CHECK(H::point_words_valid(inst), HALO_EREJECT, "instance: C is not a point of the curve");
    HALO_CHECK(H::scalar_words_valid(inst + 13) && H::scalar_words_valid(inst + 17), HALO_EREJECT, "instance: z or v not below the modulus");
    const uint64_t* pr = q.proof();
    for (unsigned i = 0; i < 2 * L + 1; ++i)               // Ls, Rs, U
        HALO_CHECK(H::point_words_valid(pr + 12 * i), HALO_EREJECT, "instance: L_i, R_i or U is not a point of the curve");
    HALO_CHECK(H::scalar_words_valid(pr + 24 * L + 12), HALO_EREJECT, "instance: c not below the modulus");
    const uint64_t hiding = pr[24 * L + 16];
    HALO_CHECK(hiding <= 1, HALO_EREJECT, "instance: malformed Option tag");
    if (hiding) {
        HALO_CHECK(H::point_words_valid(pr + 24 * L + 17), HALO_EREJECT, "instance: C_bar is not a point of the curve");
        HALO_CHECK(H::scalar_words_valid(pr + 24 * L + 29), HALO_EREJECT, "instance: w' not below the modulus");
    }
    return HALO_OK;
}

// sum_i ks[i] * pts[i] on the host for a few dozen points: 4-bit fixed windows with the doublings shared by all
// points (Straus), 252 doublings + ~74 additions per point instead of a 255-step ladder per point
template <int C>
static typename Host<C>::Pt host_msm(const typename Host<C>::Pt* pts, const typename Host<C>::SF* ks, size_t k) {
    typedef Host<C> H;
    typedef typename H::Pt Pt;
    std::vector<Pt> tab(k * 15);                            // tab[15 i + d - 1] = d * pts[i]
    std::vector<Bits256> bits(k);
    for (size_t i = 0; i < k; ++i) {
        Pt* t = tab.data() + 15 * i;
        t[0] = pts[i];
        t[1] = xyzz_dbl<C>(pts[i]);
        for (int d = 3; d <= 15; ++d) { t[d - 1] = t[d - 2]; xyzz_add<C>(t[d - 1], pts[i]); }
        bits[i] = H::bits(ks[i]);
    }
    Pt acc = xyzz_identity<C>();
    for (int w = 63; w >= 0; --w) {
        for (int r = 0; r < 4; ++r) acc = xyzz_dbl<C>(acc);
        for (size_t i = 0; i < k; ++i) {
            const u32 d = (bits[i].w[w >> 3] >> ((w & 7) * 4)) & 15u;
            if (d) xyzz_add<C>(acc, tab[15 * i + d - 1]);
        }
    }
    return acc;
}

// pcdl::succinct_check (pcdl.rs:483-554).  xis_out: (lg n + 1) elements.  HALO_EREJECT when the verifier rejects.
template <int C>
int pcdl_succinct_check(const uint64_t* inst, uint64_t* xis_out, uint64_t* U_out) {
    typedef Host<C> H;
    typedef typename H::SF SF;
    typedef typename H::Pt Pt;
    constexpr int S = H::S;
    Runtime& R = rt();
    if (int e = instance_validate<C>(inst)) return e;
    InstanceView<C> q(inst);
    const unsigned lg_n = q.lg_n;
    const SF z = q.z(), v = q.v();
    Transcript<C> tr(LABEL_PCDL);
    Pt Cp = q.Cm();
    if (q.hiding()) {
        const Pt Cbar = q.Cbar();
        tr.absorb_point(Cp); tr.absorb_point(Cbar);
        tr.absorb_scalar(z); tr.absorb_scalar(v);
        const SF a = tr.challenge();
        Cp = H::add(H::add(Cp, H::mul(Cbar, a)), H::neg(H::mul(H::from_affine_words(R.pp[C].s_xy), q.wprime())));
    }
    tr.absorb_point(Cp);
    tr.absorb_scalar(z); tr.absorb_scalar(v);
    std::vector<SF> xis(lg_n + 1);
    xis[0] = tr.challenge();
    const Pt Hp = H::mul(H::from_affine_words(R.pp[C].h_xy), xis[0]);
    // C_(lg n) = C' + v H' + sum_i (xi_i^-1 L_i + xi_i R_i)  (pcdl.rs:512-537) as ONE small multi-scalar product
    std::vector<Pt> pts(2 * lg_n + 1);
    std::vector<SF> ks(2 * lg_n + 1);
    pts[0] = Hp; ks[0] = v;
    for (unsigned i = 0; i < lg_n; ++i) {
        const Pt Li = q.Lr(i), Ri = q.Rr(i);
        tr.absorb_scalar(xis[i]);
        tr.absorb_point(Li); tr.absorb_point(Ri);
        xis[i + 1] = tr.challenge();
        HALO_CHECK(!fe_is_zero<S>(xis[i + 1]), HALO_EINVAL, "succinct_check: zero challenge has no inverse");
        pts[1 + 2 * i] = Li; ks[1 + 2 * i] = fe_inv<S>(xis[i + 1]);
        pts[2 + 2 * i] = Ri; ks[2 + 2 * i] = xis[i + 1];
    }
    const Pt Ci = H::add(Cp, host_msm<C>(pts.data(), ks.data(), pts.size()));
    const SF c = q.c();
    const SF vprime = fe_mul<S>(c, hpoly_eval<S>(xis.data(), lg_n, z));
    const Pt rhs_pts[2] = {q.U(), Hp};
    const SF rhs_ks[2] = {c, vprime};
    const Pt rhs = host_msm<C>(rhs_pts, rhs_ks, 2);
    HALO_CHECK(xyzz_eq<C>(Ci, rhs), HALO_EREJECT, "C_(log_n) != CM.Commit_Sigma(c || v')");
    if (xis_out) memcpy(xis_out, xis.data(), (lg_n + 1) * 32);
    if (U_out) memcpy(U_out, q.proof() + 24 * lg_n, 96);
    return HALO_OK;
}

// h(X) coefficients on the device: out[j] = sum_i alphas[i] * h_i[j]; xis/alphas are HOST arrays
template <int S>
static int hpoly_expand_dev(const Fe<S>* xis, const Fe<S>* alphas, size_t m, unsigned lg_n, Fe<S>* d_out) {
    Runtime& R = rt();
    Fe<S>* d_in;
    const size_t words = m * (lg_n + 1) + m;
    if (int e = ws_get(WS_TMP0, words * 32, (void**)&d_in)) return e;
    HALO_CUDA(cudaMemcpyAsync(d_in, xis, m * (lg_n + 1) * 32, cudaMemcpyHostToDevice, R.stream));
    HALO_CUDA(cudaMemcpyAsync(d_in + m * (lg_n + 1), alphas, m * 32, cudaMemcpyHostToDevice, R.stream));
    const size_t n = (size_t)1 << lg_n;
    ipa_hpoly_kernel<S><<<(unsigned)((n / 16 + 128) / 128), 128, 0, R.stream>>>(d_in, d_in + m * (lg_n + 1), m, (int)lg_n, d_out);
    R.launches += 1;
    HALO_CUDA(cudaGetLastError());
    HALO_CUDA(cudaStreamSynchronize(R.stream));               // xis/alphas may be stack memory of the caller
    return HALO_OK;
}

// commit(h) for one challenge vector: the size-n MSM of pcdl::check / acc::decider (pcdl.rs:577-580)
template <int C>
static int hpoly_commit(const typename Host<C>::SF* xis, unsigned lg_n, typename Host<C>::Pt* out) {
    typedef Host<C> H;
    constexpr int S = H::S;
    const size_t n = (size_t)1 << lg_n;
    typename H::SF* d_h;
    if (int e = ws_get(WS_IPA_C, n * 32, (void**)&d_h)) return e;
    typename H::SF one = fe_one<S>();
    if (int e = hpoly_expand_dev<S>(xis, &one, 1, lg_n, d_h)) return e;
    return commit_dev<C>(d_h, n, nullptr, out);
}

// pcdl::check (pcdl.rs:563-583) = acc::decider (acc.rs:215-217)
template <int C>
int pcdl_check(const uint64_t* inst) {
    typedef Host<C> H;
    if (int e = instance_validate<C>(inst)) return e;
    InstanceView<C> q(inst);
    std::vector<typename H::SF> xis(q.lg_n + 1);
    uint64_t U[12];
    if (int e = pcdl_succinct_check<C>(inst, (uint64_t*)xis.data(), U)) return e;
    typename H::Pt comm;
    if (int e = hpoly_commit<C>(xis.data(), q.lg_n, &comm)) return e;
    HALO_CHECK(xyzz_eq<C>(H::from_words(U), comm), HALO_EREJECT, "U != CM.Commit(ck, h_vec)");
    return HALO_OK;
}

// ---------------------------------------------------------------------------------------------------------
// AS_DL (acc.rs:128-217)
// ---------------------------------------------------------------------------------------------------------
template <int C>
struct AccCommon {
    typedef Host<C> H;
    typename H::Pt Cbar;
    size_t d;
    unsigned lg_n;
    typename H::SF z;
    std::vector<typename H::SF> xis;      // m x (lg n + 1)
    std::vector<typename H::SF> alphas;   // m
    typename H::SF eval(const typename H::SF& x) const {          // AccumulatedHPolys::eval (acc.rs:95-101)
        constexpr int S = H::S;
        typename H::SF v = fe_zero<S>();
        for (size_t i = 0; i < alphas.size(); ++i)
            v = fe_add<S>(v, fe_mul<S>(hpoly_eval<S>(xis.data() + i * (lg_n + 1), lg_n, x), alphas[i]));
        return v;
    }
};

// acc::common_subroutine (acc.rs:128-176).  insts: m flat instances of the same d.
template <int C>
static int acc_common(const uint64_t* insts, size_t m, AccCommon<C>* out, bool sharded = false) {
    typedef Host<C> H;
    typedef typename H::SF SF;
    constexpr int S = H::S;
    HALO_CHECK(m > 0 && insts, HALO_EREJECT, "No instances given");
    // the per-instance stride comes from the first instance's d: it must be a sane size BEFORE anything is read at
    // offsets derived from it (every instance is then validated in full by its succinct check)
    {
        const uint64_t d0 = insts[12];
        HALO_CHECK(d0 != ~(uint64_t)0 && ((d0 + 1) & d0) == 0, HALO_EREJECT, "acc: d + 1 of the first instance is not a power of two");
        HALO_CHECK(rt().pp[C].n > 0, HALO_EPARAMS, "acc: public parameters not loaded");
        HALO_CHECK(d0 <= rt().pp[C].n - 1, HALO_EREJECT, "d was larger than D!");
    }
    InstanceView<C> first(insts);
    const size_t d = first.d;
    const unsigned lg_n = first.lg_n;
    const size_t iw = instance_words(lg_n);
    out->d = d; out->lg_n = lg_n;
    out->xis.resize(m * (lg_n + 1));
    std::vector<uint64_t> Us(12 * m);
    const bool trace = rt().tune[7] == 1;
    auto t0 = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!trace) return;
        auto t1 = std::chrono::steady_clock::now();
        fprintf(stderr, "[halo trace] acc_common m=%zu %-24s %9.3f ms\n", m, what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    };
    // Multi-GPU (one process per GPU, SURVEY section 8(e) "ASDL over m instances: by instance"): rank r runs the succinct
    // checks of instances r, r + N, r + 2N, ... on its share of the host's cores; challenge vectors, U's and verdicts
    // are all-gathered block by block (a few hundred KB in all) while rank 0 feeds them to the serial sponge.
    if (sharded && dist_on() && dist_world() > 1) {
        // Pipelined in blocks of BS instances per rank: the ranks check block b + 1 while the all-gather of block b is
        // in flight and rank 0's sponge thread is absorbing the blocks before it -- the sponge is one serial chain of
        // m (lg n + 1) permutations and stays the critical path, as on one GPU; the other ranks spend their cores on
        // checks only and receive (alpha, z) at the end.
        const size_t W = (size_t)dist_world(), me = (size_t)dist_rank();
        const size_t BS = 16, per_blk = W * BS, nblocks = (m + per_blk - 1) / per_blk;
        const size_t rec = (size_t)(lg_n + 1) * 4 + 12 + 1;                              // words per instance record
        const size_t ntask = nblocks * BS;                                               // this rank's tasks, in order
        std::vector<uint64_t> mine(ntask * rec, 0), all(nblocks * per_blk * rec, 0);
        std::vector<std::atomic<int>> tdone(ntask);
        for (auto& f : tdone) f.store(0, std::memory_order_relaxed);
        std::atomic<size_t> nxt{0}, ready{0};                                            // ready: blocks gathered so far
        std::atomic<int> stop{0};
        auto inst_of = [&](size_t b, size_t r, size_t j) { return b * per_blk + j * W + r; };
        auto share = [&]() {
            for (;;) {
                const size_t t = nxt.fetch_add(1, std::memory_order_relaxed);
                if (t >= ntask || stop.load(std::memory_order_relaxed)) return;
                const size_t i = inst_of(t / BS, me, t % BS);
                if (i < m) {
                    const uint64_t* qi = insts + i * iw;
                    uint64_t* r = mine.data() + t * rec;
                    int code = HALO_EREJECT;
                    if ((size_t)qi[12] == d) code = pcdl_succinct_check<C>(qi, r, r + (size_t)(lg_n + 1) * 4);
                    r[rec - 1] = (uint64_t)(int64_t)code;
                }
                tdone[t].store(1, std::memory_order_release);
            }
        };
        // rank 0: the serial sponge over gathered blocks, in instance order
        Transcript<C> trs(LABEL_ASDL);
        auto rec_of = [&](size_t i) {                                                    // record of instance i in `all`
            const size_t b = i / per_blk, o = i % per_blk;
            return all.data() + ((b * W + o % W) * BS + o / W) * rec;
        };
        auto sponge = [&]() {
            for (size_t b = 0; b < nblocks; ++b) {
                while (ready.load(std::memory_order_acquire) <= b) { if (stop.load(std::memory_order_relaxed)) return; std::this_thread::yield(); }
                for (size_t i = b * per_blk; i < std::min(m, (b + 1) * per_blk); ++i) {
                    const Fe<S>* x = (const Fe<S>*)rec_of(i);
                    for (unsigned k = 0; k <= lg_n; ++k) trs.absorb_scalar(x[k]);
                }
            }
        };
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        const size_t cores = std::max<size_t>(1, hw / W);
        Workers wk, sp;
        if (me == 0) sp.spawn(1, sponge);
        const bool own_sponge_inline = me == 0 && sp.size() == 0;
        wk.spawn((unsigned)std::max<size_t>(1, cores > (me == 0 ? 2u : 1u) ? cores - (me == 0 ? 2 : 1) : 1), share);
        int bad_code = 0;
        int gather_rc = HALO_OK;
        for (size_t b = 0; b < nblocks && !bad_code && !gather_rc; ++b) {
            for (size_t j = 0; j < BS; ++j)
                while (!tdone[b * BS + j].load(std::memory_order_acquire)) { if (wk.size() == 0) share(); else std::this_thread::yield(); }
            gather_rc = dist_allgather_host(mine.data() + b * BS * rec, all.data() + b * per_blk * rec, BS * rec * 8);
            if (gather_rc) break;
            for (size_t i = b * per_blk; i < std::min(m, (b + 1) * per_blk) && !bad_code; ++i)
                bad_code = (int)(int64_t)rec_of(i)[rec - 1];
            if (!bad_code) ready.store(b + 1, std::memory_order_release);
        }
        if (bad_code || gather_rc) stop.store(1, std::memory_order_relaxed);
        wk.join();
        if (own_sponge_inline && !bad_code && !gather_rc) sponge();
        sp.join();
        if (gather_rc) return gather_rc;
        if (bad_code) { last_err() = "acc: an instance failed its succinct check (sharded)"; return bad_code; }
        for (size_t i = 0; i < m; ++i) {
            const uint64_t* r = rec_of(i);
            memcpy(out->xis.data() + i * (lg_n + 1), r, (size_t)(lg_n + 1) * 32);
            memcpy(Us.data() + 12 * i, r + (size_t)(lg_n + 1) * 4, 96);
        }
        lap("sharded succinct checks + gathers + sponge");
        // alpha and z come from rank 0's transcript (z is the next squeeze after alpha, as in the single-GPU path below)
        SF az[2] = {fe_zero<S>(), fe_zero<S>()};
        if (me == 0) {
            for (size_t i = 0; i < m; ++i) trs.absorb_point(H::from_words(Us.data() + 12 * i));
            az[0] = trs.challenge();
            az[1] = trs.challenge();
        }
        std::vector<SF> gath(2 * W);
        if (int e = dist_allgather_host(az, gath.data(), 64)) return e;
        const SF alpha = gath[0];
        out->z = gath[1];
        out->alphas.resize(m);
        SF cur = fe_one<S>();
        for (size_t i = 0; i < m; ++i) { out->alphas[i] = cur; cur = fe_mul<S>(cur, alpha); }
        uint64_t Cw[12];
        if (int e = halo_msm_proj(C, Us.data(), (const uint64_t*)out->alphas.data(), m, Cw)) return e;
        out->Cbar = H::from_words(Cw);
        return HALO_OK;
    }
    // 4. m independent succinct checks (each sequential inside) on worker threads that take the instances in
    // order, while this thread already feeds finished challenge vectors to the AS_DL sponge (step 6 is one serial
    // chain of ~m lg n permutations - as long as all the checks together at m = 1024)
    std::vector<int> rc(m, HALO_OK);
    std::vector<std::string> why(m);
    std::vector<std::atomic<int>> done(m);
    for (auto& f : done) f.store(0, std::memory_order_relaxed);
    std::atomic<size_t> next{0};
    auto work = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1, std::memory_order_relaxed);
            if (i >= m) return;
            const uint64_t* qi = insts + i * iw;
            if ((size_t)qi[12] != d) { rc[i] = HALO_EREJECT; why[i] = "d_i != d"; }
            else {
                rc[i] = pcdl_succinct_check<C>(qi, (uint64_t*)(out->xis.data() + i * (lg_n + 1)), Us.data() + 12 * i);
                if (rc[i]) why[i] = last_err();
            }
            done[i].store(1, std::memory_order_release);
        }
    };
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nt = (unsigned)std::min<size_t>(m, hw > 1 ? hw - 1 : 1);
    Workers th;
    if (m > 1) th.spawn(nt, work);
    if (th.size() == 0) work();
    // 6. alpha = rho1([h_i, U_i])
    Transcript<C> tr(LABEL_ASDL);
    int first_bad = -1;
    for (size_t i = 0; i < m && first_bad < 0; ++i) {
        while (!done[i].load(std::memory_order_acquire)) std::this_thread::yield();
        if (rc[i]) { first_bad = (int)i; break; }
        for (unsigned k = 0; k <= lg_n; ++k) tr.absorb_scalar(out->xis[i * (lg_n + 1) + k]);
    }
    th.join();
    if (first_bad < 0) for (size_t i = 0; i < m; ++i) if (rc[i]) { first_bad = (int)i; break; }
    if (first_bad >= 0) { last_err() = why[first_bad]; return rc[first_bad]; }
    lap("succinct checks + sponge");
    for (size_t i = 0; i < m; ++i) tr.absorb_point(H::from_words(Us.data() + 12 * i));
    const SF alpha = tr.challenge();
    lap("transcript");
    out->alphas.resize(m);
    SF cur = fe_one<S>();
    for (size_t i = 0; i < m; ++i) { out->alphas[i] = cur; cur = fe_mul<S>(cur, alpha); }
    // 8. C = sum alpha^i U_i  (point_dot, group.rs:53-56)
    uint64_t Cw[12];
    if (int e = halo_msm_proj(C, Us.data(), (const uint64_t*)out->alphas.data(), m, Cw)) return e;
    out->Cbar = H::from_words(Cw);
    // 9. z = rho1(C, h)
    out->z = tr.challenge();
    lap("C_bar, z");
    return HALO_OK;
}

// acc::prover (acc.rs:178-198): out_acc is a flat instance
template <int C>
int acc_prover(const uint64_t* insts, size_t m, uint64_t* out_acc, bool shard_open) {
    typedef Host<C> H;
    typedef typename H::SF SF;
    constexpr int S = H::S;
    AccCommon<C> cs;
    if (int e = acc_common<C>(insts, m, &cs)) return e;
    const SF v = cs.eval(cs.z);
    const size_t n = cs.d + 1;
    // h(X) = sum alpha^i h_i(X), expanded straight into the opening's coefficient workspace
    SF* d_h;
    if (int e = ws_get(WS_IPA_C, n * 32, (void**)&d_h)) return e;
    if (int e = hpoly_expand_dev<S>(cs.xis.data(), cs.alphas.data(), m, cs.lg_n, d_h)) return e;
    H::to_words(cs.Cbar, out_acc);
    out_acc[12] = cs.d;
    H::put_scalar(cs.z, out_acc + 13);
    H::put_scalar(v, out_acc + 17);
    uint64_t Cw[12];
    H::to_words(cs.Cbar, Cw);
    // pcdl::open evaluates h(z) itself (pcdl.rs:471); numerically the same value as v
    return pcdl_open<C>(nullptr, n, Cw, cs.d, (const uint64_t*)&cs.z, nullptr, nullptr, nullptr, nullptr, out_acc + 21,
                        nullptr, true, shard_open);
}

// sum of `parts` vectors of n elements: out[i] = sum_p in[p * n + i]  (modular addition is not an NCCL operator)
template <int F>
__global__ void __launch_bounds__(256) vec_sum_kernel(const Fe<F>* __restrict__ in, size_t n, int parts, Fe<F>* __restrict__ out) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fe<F> s = g_ld(in + i);
    for (int p = 1; p < parts; ++p) s = fe_add<F>(s, g_ld(in + (size_t)p * n + i));
    g_st(out + i, s);
}

// acc::prover across the GPUs of the communicator: succinct checks by instance (acc_common), h(X) = sum alpha^i h_i(X)
// as per-rank partial coefficient vectors over an instance range, all-gathered and added on every rank; the opening's
// rounds above the fold-free tail are sharded too (pcdl_open with shard = true); every rank returns the same accumulator.
template <int C>
int acc_prover_dist(const uint64_t* insts, size_t m, uint64_t* out_acc) {
    typedef Host<C> H;
    typedef typename H::SF SF;
    constexpr int S = H::S;
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    AccCommon<C> cs;
    if (int e = acc_common<C>(insts, m, &cs, true)) return e;
    const SF v = cs.eval(cs.z);
    const size_t n = cs.d + 1;
    const int W = dist_world();
    SF *d_h, *d_part;
    if (int e = ws_get(WS_IPA_C, n * 32, (void**)&d_h)) return e;
    if (int e = ws_get(WS_IPA_T1, (size_t)(W + 1) * n * 32, (void**)&d_part)) return e;
    size_t lo, hi;
    dist_range(m, &lo, &hi);
    if (hi > lo) {
        if (int e = hpoly_expand_dev<S>(cs.xis.data() + lo * (cs.lg_n + 1), cs.alphas.data() + lo, hi - lo, cs.lg_n, d_part)) return e;
    } else {
        HALO_CUDA(cudaMemsetAsync(d_part, 0, n * 32, rt().stream));
    }
    if (int e = dist_allgather_dev(d_part, d_part + n, n * 32)) return e;
    vec_sum_kernel<S><<<(unsigned)((n + 255) / 256), 256, 0, rt().stream>>>(d_part + n, n, W, d_h);
    rt().launches += 1;
    HALO_CUDA(cudaGetLastError());
    H::to_words(cs.Cbar, out_acc);
    out_acc[12] = cs.d;
    H::put_scalar(cs.z, out_acc + 13);
    H::put_scalar(v, out_acc + 17);
    uint64_t Cw[12];
    H::to_words(cs.Cbar, Cw);
    return pcdl_open<C>(nullptr, n, Cw, cs.d, (const uint64_t*)&cs.z, nullptr, nullptr, nullptr, nullptr, out_acc + 21,
                        nullptr, true, true);
}
template int acc_prover_dist<0>(const uint64_t*, size_t, uint64_t*);
template int acc_prover_dist<1>(const uint64_t*, size_t, uint64_t*);

// acc::decider / pcdl::check with commit(h) as a point-range-sharded MSM over the resident generators
template <int C>
int pcdl_check_dist(const uint64_t* inst) {
    typedef Host<C> H;
    constexpr int S = H::S;
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    if (int e = instance_validate<C>(inst)) return e;
    InstanceView<C> q(inst);
    std::vector<typename H::SF> xis(q.lg_n + 1);
    uint64_t U[12];
    if (int e = pcdl_succinct_check<C>(inst, (uint64_t*)xis.data(), U)) return e;
    const size_t n = (size_t)1 << q.lg_n;
    typename H::SF* d_h;
    if (int e = ws_get(WS_IPA_C, n * 32, (void**)&d_h)) return e;
    typename H::SF one = fe_one<S>();
    if (int e = hpoly_expand_dev<S>(xis.data(), &one, 1, q.lg_n, d_h)) return e;
    size_t lo, hi;
    dist_range(n, &lo, &hi);
    uint64_t cw[12];
    if (int e = dist_msm_device<C>((const Affine<C>*)rt().pp[C].d_gs + lo, d_h + lo, hi - lo, 1, cw)) return e;
    HALO_CHECK(xyzz_eq<C>(H::from_words(U), H::from_words(cw)), HALO_EREJECT, "U != CM.Commit(ck, h_vec)");
    return HALO_OK;
}
template int pcdl_check_dist<0>(const uint64_t*);
template int pcdl_check_dist<1>(const uint64_t*);

// acc::verifier (acc.rs:200-213)
template <int C>
int acc_verifier(const uint64_t* insts, size_t m, const uint64_t* acc, bool sharded) {
    typedef Host<C> H;
    constexpr int S = H::S;
    AccCommon<C> cs;
    if (int e = acc_common<C>(insts, m, &cs, sharded)) return e;
    {
        const int e = instance_validate<C>(acc);
        if (e) return e == HALO_EINVAL ? HALO_EREJECT : e;
    }
    InstanceView<C> a(acc);
    HALO_CHECK(xyzz_eq<C>(cs.Cbar, a.Cm()), HALO_EREJECT, "C_bar' != C_bar");
    HALO_CHECK(fe_eq<S>(cs.z, a.z()), HALO_EREJECT, "z' = z");
    HALO_CHECK(cs.d == a.d, HALO_EREJECT, "d' = d");
    HALO_CHECK(fe_eq<S>(cs.eval(a.z()), a.v()), HALO_EREJECT, "h(z) = v");
    return HALO_OK;
}

template int pcdl_commit<0>(const uint64_t*, size_t, size_t, const uint64_t*, uint64_t*);
template int pcdl_commit<1>(const uint64_t*, size_t, size_t, const uint64_t*, uint64_t*);
template int pcdl_open<0>(const uint64_t*, size_t, const uint64_t*, size_t, const uint64_t*, const uint64_t*, const uint64_t*,
                          const uint64_t*, const uint64_t*, uint64_t*, uint64_t*, bool, bool);
template int pcdl_open<1>(const uint64_t*, size_t, const uint64_t*, size_t, const uint64_t*, const uint64_t*, const uint64_t*,
                          const uint64_t*, const uint64_t*, uint64_t*, uint64_t*, bool, bool);
template int pcdl_succinct_check<0>(const uint64_t*, uint64_t*, uint64_t*);
template int pcdl_succinct_check<1>(const uint64_t*, uint64_t*, uint64_t*);
template int pcdl_check<0>(const uint64_t*);
template int pcdl_check<1>(const uint64_t*);
template int acc_prover<0>(const uint64_t*, size_t, uint64_t*, bool);
template int acc_prover<1>(const uint64_t*, size_t, uint64_t*, bool);
template int acc_verifier<0>(const uint64_t*, size_t, const uint64_t*, bool);
template int acc_verifier<1>(const uint64_t*, size_t, const uint64_t*, bool);

// h(X) coefficient vector for m accumulated instances, returned to the host (≙ AccumulatedHPolys::get_poly)
template <int S>
int hpoly_coeffs(const uint64_t* xis, const uint64_t* alphas, size_t m, unsigned lg_n, uint64_t* out) {
    Runtime& R = rt();
    const size_t n = (size_t)1 << lg_n;
    Fe<S>* d_h;
    if (int e = ws_get(WS_IPA_C, n * 32, (void**)&d_h)) return e;
    if (int e = hpoly_expand_dev<S>((const Fe<S>*)xis, (const Fe<S>*)alphas, m, lg_n, d_h)) return e;
    HALO_CUDA(cudaMemcpyAsync(out, d_h, n * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
}
template int hpoly_coeffs<0>(const uint64_t*, const uint64_t*, size_t, unsigned, uint64_t*);
template int hpoly_coeffs<1>(const uint64_t*, const uint64_t*, size_t, unsigned, uint64_t*);

template <int C>
int hpoly_commit_words(const uint64_t* xis, unsigned lg_n, uint64_t* out_xyz) {
    typename Host<C>::Pt p;
    if (int e = hpoly_commit<C>((const typename Host<C>::SF*)xis, lg_n, &p)) return e;
    Host<C>::to_words(p, out_xyz);
    return HALO_OK;
}
template int hpoly_commit_words<0>(const uint64_t*, unsigned, uint64_t*);
template int hpoly_commit_words<1>(const uint64_t*, unsigned, uint64_t*);

template <int S>
void hpoly_eval_words(const uint64_t* xis, unsigned lg_n, const uint64_t* z, uint64_t* out) {
    Fe<S> zz; memcpy(&zz, z, 32);
    Fe<S> v = hpoly_eval<S>((const Fe<S>*)xis, lg_n, zz);
    memcpy(out, &v, 32);
}
template void hpoly_eval_words<0>(const uint64_t*, unsigned, const uint64_t*, uint64_t*);
template void hpoly_eval_words<1>(const uint64_t*, unsigned, const uint64_t*, uint64_t*);


// ---- group.rs:43-66 exposed directly: scalar_dot and construct_powers on the device, host in / host out ----
template <int S>
int scalar_dot_host(const uint64_t* a, const uint64_t* b, size_t n, uint64_t* out) {
    Runtime& R = rt();
    Fe<S> acc = fe_zero<S>();
    if (n) {
        Fe<S>*da, *db, *part;
        if (int e = ws_get(WS_IPA_C, n * 32, (void**)&da)) return e;
        if (int e = ws_get(WS_IPA_Z, n * 32, (void**)&db)) return e;
        if (int e = ws_get(WS_IPA_T0, 2 * 256 * 32 + 64, (void**)&part)) return e;
        HALO_CUDA(cudaMemcpyAsync(da, a, n * 32, cudaMemcpyHostToDevice, R.stream));
        HALO_CUDA(cudaMemcpyAsync(db, b, n * 32, cudaMemcpyHostToDevice, R.stream));
        const int g = (int)std::min<size_t>(256, (n + 255) / 256);
        ipa_dot1_kernel<S><<<g, 256, 0, R.stream>>>(da, db, n, part);
        R.launches += 1;
        if (int e = sum_partials<S>(part, g, &acc, 1)) return e;
    }
    memcpy(out, &acc, 32);
    return HALO_OK;
}
template int scalar_dot_host<0>(const uint64_t*, const uint64_t*, size_t, uint64_t*);
template int scalar_dot_host<1>(const uint64_t*, const uint64_t*, size_t, uint64_t*);

template <int S>
int construct_powers_host(const uint64_t* z, size_t n, uint64_t* out) {
    Runtime& R = rt();
    if (n == 0) return HALO_OK;
    Fe<S>* d;
    if (int e = ws_get(WS_IPA_Z, n * 32, (void**)&d)) return e;
    Fe<S> zz; memcpy(&zz, z, 32);
    ipa_powers_kernel<S><<<(unsigned)((n / 16 + 128) / 128), 128, 0, R.stream>>>(d, n, zz, fe_one<S>());
    R.launches += 1;
    HALO_CUDA(cudaGetLastError());
    HALO_CUDA(cudaMemcpyAsync(out, d, n * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
}
template int construct_powers_host<0>(const uint64_t*, size_t, uint64_t*);
template int construct_powers_host<1>(const uint64_t*, size_t, uint64_t*);

}  // namespace halo
