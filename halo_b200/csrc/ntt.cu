// Radix-2 NTT / iNTT over the two Pasta fields on sm_100a.
//
// Replaces ark-poly 0.5.0's `GeneralEvaluationDomain` transforms as reached from
//   crates/group/src/poly.rs:56-64 (Evals::from_poly*  -> evaluate_over_domain_by_ref)
//   crates/group/src/poly.rs:133-139 (Evals::interpolate*)
//   crates/plonk/src/plonk/protocol.rs:82-106,140-141,158-159,195 and crates/plonk/src/circuit/trace.rs:171-177.
// Conventions: omega_N = (5^((p-1)/2^32))^(2^(32-log N)); forward
// evals[i] = sum_j coeff[j] omega^(ij); natural order in and out; the inverse includes the factor N^-1.  The
// coset variants (evaluate on g*<omega>) are an API addition: the reference has no coset call site.
//
// Schedule: decimation in frequency.  Stage t pairs elements 2^(n-1-t) apart:
//     a' = a + b,   b' = (a - b) * omega_N^((i mod 2^(n-1-t)) << t)
// and after n stages element i holds X[bitrev(i)].  The n stages are cut into passes; a pass loads a tile of
// 2^(s+q) elements into shared memory (2^s strided "rows" of 2^q consecutive elements, so every global access is a
// contiguous 32*2^q-byte run), runs its s stages there, and writes the tile back in place.  The last pass owns the
// lowest stages (contiguous tiles) and writes each element to its bit-reversed position in the output buffer,
// multiplied by N^-1 for the inverse.  Twiddles come from one table per field, T[k] = omega_max^k, k < 2^(logmax-1):
// smaller domains read it with a stride, the inverse transform reads omega^-k = -omega^(N/2-k).
// Work: (N/2) log2 N butterflies = one Montgomery multiplication + one add + one sub each; 64 N bytes of HBM
// traffic per pass.  At 32 B per element this is integer-pipe-bound, not HBM-bound.
#include "runtime.cuh"
#include <algorithm>

namespace halo {

constexpr int NTT_LOG_TILE = 11;             // 2048 elements = 64 KB of shared memory per CTA
constexpr int NTT_THREADS = 256;
constexpr int NTT_Q = 2;                     // strided passes move runs of at least 4 elements = 128 B (q = 11 - s)

template <int F> __device__ __forceinline__ Fe<F> ldg_fe(const Fe<F>* p) {
    Fe<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int F> __device__ __forceinline__ Fe<F> ld_fe(const Fe<F>* p) {
    Fe<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int F> __device__ __forceinline__ void st_fe(Fe<F>* p, const Fe<F>& v) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    q[1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}
// shared-memory tile as two uint4 planes (low / high half of every element): consecutive elements are 16 B
// apart in each plane, so a quarter-warp's LDS.128 covers 128 contiguous bytes (no bank conflicts)
template <int F> __device__ __forceinline__ Fe<F> sm_get(const uint4* lo, const uint4* hi, u32 e) {
    Fe<F> r;
    uint4 a = lo[e], b = hi[e];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int F> __device__ __forceinline__ void sm_put(uint4* lo, uint4* hi, u32 e, const Fe<F>& v) {
    lo[e] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    hi[e] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}

struct NttPass {
    int n;          // log2 N
    int t0, s;      // first stage and number of stages of this pass
    int q;          // log2 of the contiguous run length (0 for the last pass)
    int tw_shift;   // logmax - n: stride of this domain's twiddles in the table
    int inverse;
    int last;       // write bit-reversed (and scaled) to `out`
};

// one CTA per tile, blockIdx.y = polynomial of the batch
template <int F>
__global__ void __launch_bounds__(NTT_THREADS) ntt_pass_kernel(Fe<F>* __restrict__ data, Fe<F>* __restrict__ out,
                                                               const Fe<F>* __restrict__ tw, NttPass P, Fe<F> scale) {
    extern __shared__ uint4 smem[];
    const u32 tile = 1u << (P.s + P.q);
    uint4* lo = smem;
    uint4* hi = smem + tile;
    const int lowbits = P.n - P.t0 - P.s;
    const u32 T = blockIdx.x;
    const u32 lp = T & ((1u << (lowbits - P.q)) - 1u);
    const u32 high = T >> (lowbits - P.q);
    const size_t N = (size_t)1 << P.n;
    Fe<F>* x = data + (size_t)blockIdx.y * N;
    const size_t base = ((size_t)high << (P.n - P.t0)) | ((size_t)lp << P.q);
    const u32 qmask = (1u << P.q) - 1u;

    for (u32 e = threadIdx.x; e < tile; e += NTT_THREADS) {
        const size_t i = base | ((size_t)(e >> P.q) << lowbits) | (e & qmask);
        sm_put<F>(lo, hi, e, ld_fe(x + i));
    }
    __syncthreads();

    const u32 half_n = 1u << (P.n - 1);
    for (int st = 0; st < P.s; ++st) {
        const int t = P.t0 + st;
        const int u = P.s - 1 - st;                       // bit of the row index paired by this stage
        const u32 umask = (1u << u) - 1u;
        if (t == P.n - 1) {
            // the last stage of the transform: j = 0 for every butterfly, the twiddle is omega^0 = 1 in both directions,
            // and a product with one returns its (already reduced) operand -- a twentieth of a 2^20 transform's
            // multiplications that need not be done
            for (u32 b = threadIdx.x; b < (tile >> 1); b += NTT_THREADS) {
                const u32 l = b & qmask, r = b >> P.q;
                const u32 m0 = ((r >> u) << (u + 1)) | (r & umask);
                const u32 e0 = (m0 << P.q) | l, e1 = e0 + (1u << (u + P.q));
                Fe<F> a = sm_get<F>(lo, hi, e0), c = sm_get<F>(lo, hi, e1);
                sm_put<F>(lo, hi, e0, fe_add<F>(a, c));
                sm_put<F>(lo, hi, e1, fe_sub<F>(a, c));
            }
            __syncthreads();
            continue;
        }
        for (u32 b = threadIdx.x; b < (tile >> 1); b += NTT_THREADS) {
            const u32 l = b & qmask, r = b >> P.q;
            const u32 m0 = ((r >> u) << (u + 1)) | (r & umask);
            const u32 e0 = (m0 << P.q) | l, e1 = e0 + (1u << (u + P.q));
            // j = i0 mod 2^(n-1-t); twiddle exponent k = j << t  (< N/2)
            const u32 j = ((m0 & umask) << lowbits) | (lp << P.q) | l;
            const u32 k = j << t;
            Fe<F> a = sm_get<F>(lo, hi, e0), c = sm_get<F>(lo, hi, e1);
            Fe<F> sum = fe_add<F>(a, c), d, w;
            if (!P.inverse) {
                d = fe_sub<F>(a, c);
                w = ldg_fe(tw + ((size_t)k << P.tw_shift));
            } else if (k == 0) {
                d = fe_sub<F>(a, c);
                w = ldg_fe(tw);
            } else {                                        // omega^-k = -omega^(N/2-k)
                d = fe_sub<F>(c, a);
                w = ldg_fe(tw + ((size_t)(half_n - k) << P.tw_shift));
            }
            sm_put<F>(lo, hi, e0, sum);
            sm_put<F>(lo, hi, e1, fe_mul<F>(d, w));
        }
        __syncthreads();
    }

    if (!P.last) {
        for (u32 e = threadIdx.x; e < tile; e += NTT_THREADS) {
            const size_t i = base | ((size_t)(e >> P.q) << lowbits) | (e & qmask);
            st_fe(x + i, sm_get<F>(lo, hi, e));
        }
    } else {
        // last pass: lowbits == 0, q == 0, i = high << s | e; natural-order position = bitrev_n(i)
        Fe<F>* y = out + (size_t)blockIdx.y * N;
        for (u32 e = threadIdx.x; e < tile; e += NTT_THREADS) {
            const size_t i = base | e;
            const size_t o = (size_t)(__brev((u32)i) >> (32 - P.n));
            Fe<F> v = sm_get<F>(lo, hi, e);
            if (P.inverse) v = fe_mul<F>(v, scale);
            st_fe(y + o, v);
        }
    }
}

// T[k] = omega^k for k < count; one thread per run of 64 consecutive powers
template <int F>
__global__ void __launch_bounds__(128) ntt_twiddle_kernel(Fe<F>* __restrict__ T, size_t count, Fe<F> omega) {
    size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t lo = c * 64;
    if (lo >= count) return;
    u32 e[8] = {(u32)lo, (u32)(lo >> 32), 0, 0, 0, 0, 0, 0};
    Fe<F> w = fe_pow<F>(omega, e, 40);
    size_t hi = lo + 64 < count ? lo + 64 : count;
    for (size_t k = lo; k < hi; ++k) { st_fe(T + k, w); w = fe_mul<F>(w, omega); }
}

// Coset scaling x[j][k] *= c0 * g^k for every vector j of a batch (pre-scaling of the forward coset transform;
// g^-k * N^-1 after the inverse).  g^k = lo[k & 1023] * hi[k >> 10] from two small tables (lo carries c0), one thread
// per k so that every access is coalesced and the power is shared by the whole batch: 1 + 1/batch multiplications
// per element.
template <int F>
__global__ void __launch_bounds__(128) ntt_power_table_kernel(Fe<F>* __restrict__ tab, size_t n, Fe<F> x, Fe<F> c0) {
    size_t lo = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 16;
    if (lo >= n) return;
    Fe<F> cur = c0, b = x;
    for (size_t e = lo; e; e >>= 1) { if (e & 1) cur = fe_mul<F>(cur, b); b = fe_sqr<F>(b); }
    for (size_t k = lo; k < lo + 16 && k < n; ++k) { st_fe(tab + k, cur); cur = fe_mul<F>(cur, x); }
}
template <int F>
__global__ void __launch_bounds__(256) ntt_coset_scale_kernel(Fe<F>* __restrict__ data, size_t N, size_t batch,
                                                              const Fe<F>* __restrict__ lo, const Fe<F>* __restrict__ hi) {
    size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (k >= N) return;
    Fe<F> w = ld_fe(lo + (k & 1023));
    if (k >> 10) w = fe_mul<F>(w, ld_fe(hi + (k >> 10)));
    for (size_t j = 0; j < batch; ++j) { Fe<F>* x = data + j * N + k; st_fe(x, fe_mul<F>(ld_fe(x), w)); }
}
template <int F>
static int coset_scale(Fe<F>* data, size_t N, size_t batch, const Fe<F>& g, const Fe<F>& c0) {
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    Fe<F>* tab;
    const size_t nhi = (N + 1023) / 1024;
    if (int e = ws_get(WS_NTT_TAB, (1024 + nhi) * sizeof(Fe<F>), (void**)&tab)) return e;
    Fe<F> g1024 = g;
    for (int i = 0; i < 10; ++i) g1024 = fe_sqr<F>(g1024);
    ntt_power_table_kernel<F><<<1, 128, 0, st>>>(tab, 1024, g, c0);
    ntt_power_table_kernel<F><<<(unsigned)((nhi / 16 + 128) / 128), 128, 0, st>>>(tab + 1024, nhi, g1024, fe_one<F>());
    ntt_coset_scale_kernel<F><<<(unsigned)((N + 255) / 256), 256, 0, st>>>(data, N, batch, tab, tab + 1024);
    R.launches += 3;
    return HALO_OK;
}

// out[i] = sum_k coeffs[i + k*N] for i < N (zero-padded when len < N): what evaluate_over_domain_by_ref does with a
// coefficient vector longer than the domain (protocol.rs:140-141) or shorter (everywhere else)
template <int F>
__global__ void __launch_bounds__(256) ntt_fold_kernel(const Fe<F>* __restrict__ coeffs, size_t len, size_t N,
                                                       Fe<F>* __restrict__ out) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= N) return;
    Fe<F> acc = fe_zero<F>();
    for (size_t k = i; k < len; k += N) acc = fe_add<F>(acc, ld_fe(coeffs + k));
    st_fe(out + i, acc);
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
template <int F> static Fe<F> host_root_of_unity(int log_n) {
    // 5^((p-1)/2^32), then squared down to order 2^log_n  (ark-ff GENERATOR = 5, TWO_ADICITY = 32)
    u32 e[8];
    for (int i = 0; i < 7; ++i) e[i] = modulus_limb<F>(i + 1);   // (p - 1) >> 32  (p0 = 1)
    e[7] = 0;
    Fe<F> w = fe_pow<F>(fe_from_u64<F>(5), e, 256);
    for (int i = log_n; i < 32; ++i) w = fe_sqr<F>(w);
    return w;
}

template <int F> static int ensure_twiddles(int log_n) {
    Runtime& R = rt();
    if (R.ntt_tw[F] && R.ntt_logmax[F] >= log_n) return HALO_OK;
    int logmax = std::max(log_n, 16);
    if (R.ntt_tw[F]) { HALO_CUDA(cudaStreamSynchronize(R.stream)); HALO_CUDA(cudaFree(R.ntt_tw[F])); R.ntt_tw[F] = nullptr; }
    size_t count = (size_t)1 << (logmax - 1);
    HALO_CUDA(cudaMalloc(&R.ntt_tw[F], count * sizeof(Fe<F>)));
    size_t chunks = (count + 63) / 64;
    ntt_twiddle_kernel<F><<<(unsigned)((chunks + 127) / 128), 128, 0, R.stream>>>((Fe<F>*)R.ntt_tw[F], count,
                                                                                   host_root_of_unity<F>(logmax));
    HALO_CUDA(cudaGetLastError());
    R.launches += 1;
    R.ntt_logmax[F] = logmax;
    static bool attr_set[2] = {false, false};
    if (!attr_set[F]) {
        HALO_CUDA(cudaFuncSetAttribute(ntt_pass_kernel<F>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (1 << NTT_LOG_TILE) * 32));
        attr_set[F] = true;
    }
    return HALO_OK;
}

// Transforms `batch` vectors of 2^log_n elements in `data`; the result is left in `out` (a distinct buffer of the
// same size).  coset == nullptr for the plain subgroup domain.
template <int F>
int ntt_device(Fe<F>* data, Fe<F>* out, int log_n, size_t batch, bool inverse, const Fe<F>* coset) {
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    const size_t N = (size_t)1 << log_n;
    if (batch == 0) return HALO_OK;
    HALO_CHECK(log_n >= 0 && log_n <= 30, HALO_EINVAL, "ntt: log_n must be in [0, 30]");
    HALO_CHECK(batch <= 65535, HALO_EINVAL, "ntt: batch must be <= 65535");
    if (log_n == 0) {                                   // N = 1: the transform is the identity in both directions
        HALO_CUDA(cudaMemcpyAsync(out, data, batch * sizeof(Fe<F>), cudaMemcpyDeviceToDevice, st));
        return HALO_OK;
    }
    if (int e = ensure_twiddles<F>(log_n)) return e;
    if (int e = prof_begin()) return e;
    const Fe<F>* tw = (const Fe<F>*)R.ntt_tw[F];
    if (coset && !inverse) {
        if (int e = coset_scale<F>(data, N, batch, *coset, fe_one<F>())) return e;
        prof_mark("coset");
    }
    Fe<F> ninv = fe_inv<F>(fe_from_u64<F>((u64)N));
    Fe<F> scale = (inverse && !coset) ? ninv : fe_one<F>();
    NttPass P;
    P.n = log_n; P.tw_shift = R.ntt_logmax[F] - log_n; P.inverse = inverse ? 1 : 0;
    const int last_s = std::min(log_n, NTT_LOG_TILE);
    const int rem = log_n - last_s;
    const int smax = NTT_LOG_TILE - NTT_Q;
    const int np = (rem + smax - 1) / smax;
    int t0 = 0;
    for (int p = 0; p < np; ++p) {
        int s = rem / np + (p < rem % np ? 1 : 0);
        // every strided pass fills the whole 2^11-element tile: 2^s rows of 2^q consecutive elements with q = 11 - s >= 2
        // (round 1 fixed q = 2, so the 5- and 6-stage passes of 2^16 / 2^22 transforms ran on tiles of 128 - 256 elements
        // with most of the CTA idle); the bits below this pass's stages number n - t0 - s >= 11 > q
        P.t0 = t0; P.s = s; P.q = NTT_LOG_TILE - s; P.last = 0;
        dim3 grid((unsigned)(N >> (s + P.q)), (unsigned)batch);
        ntt_pass_kernel<F><<<grid, NTT_THREADS, (size_t)32 << (s + P.q), st>>>(data, out, tw, P, scale);
        R.launches += 1;
        prof_mark("pass");
        t0 += s;
    }
    P.t0 = t0; P.s = last_s; P.q = 0; P.last = 1;
    {
        dim3 grid((unsigned)(N >> last_s), (unsigned)batch);
        ntt_pass_kernel<F><<<grid, NTT_THREADS, (size_t)32 << last_s, st>>>(data, out, tw, P, (inverse && coset) ? fe_one<F>() : scale);
        R.launches += 1;
        prof_mark("last_pass");
    }
    if (coset && inverse) {
        if (int e = coset_scale<F>(out, N, batch, fe_inv<F>(*coset), ninv)) return e;
        prof_mark("coset");
    }
    HALO_CUDA(cudaGetLastError());
    return HALO_OK;
}
template int ntt_device<0>(Fe<0>*, Fe<0>*, int, size_t, bool, const Fe<0>*);
template int ntt_device<1>(Fe<1>*, Fe<1>*, int, size_t, bool, const Fe<1>*);

template <int F>
int ntt_fold_device(const Fe<F>* coeffs, size_t len, int log_n, Fe<F>* out) {
    Runtime& R = rt();
    const size_t N = (size_t)1 << log_n;
    ntt_fold_kernel<F><<<(unsigned)((N + 255) / 256), 256, 0, R.stream>>>(coeffs, len, N, out);
    HALO_CUDA(cudaGetLastError());
    R.launches += 1;
    return HALO_OK;
}
template int ntt_fold_device<0>(const Fe<0>*, size_t, int, Fe<0>*);
template int ntt_fold_device<1>(const Fe<1>*, size_t, int, Fe<1>*);


// pointwise product of two evaluation vectors (the middle step of an FFT-based polynomial product)
template <int F>
__global__ void __launch_bounds__(256) ntt_pointwise_mul_kernel(Fe<F>* __restrict__ a, const Fe<F>* __restrict__ b, size_t n) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_fe(a + i, fe_mul<F>(ld_fe(a + i), ld_fe(b + i)));
}

// Product of two coefficient vectors as ark-poly's `&a * &b` computes it for DensePolynomial (protocol.rs:132-139,
// 198-199; pcdl.rs:215): evaluate both on the domain of size next_pow2(la + lb - 1), multiply pointwise, interpolate.
// out receives la + lb - 1 coefficients.  Everything stays on the device between the three transforms.
template <int F>
int poly_mul_host(const uint64_t* a, size_t la, const uint64_t* b, size_t lb, uint64_t* out) {
    Runtime& R = rt();
    if (la == 0 || lb == 0) return HALO_OK;
    const size_t lo = la + lb - 1;
    int log_n = 0; while (((size_t)1 << log_n) < lo) ++log_n;
    const size_t N = (size_t)1 << log_n;
    Fe<F>*buf, *scratch;
    if (int e = ws_get(WS_NTT0, 2 * N * 32, (void**)&buf)) return e;
    if (int e = ws_get(WS_NTT1, 2 * N * 32, (void**)&scratch)) return e;
    HALO_CUDA(cudaMemsetAsync(buf, 0, 2 * N * 32, R.stream));
    HALO_CUDA(cudaMemcpyAsync(buf, a, la * 32, cudaMemcpyHostToDevice, R.stream));
    HALO_CUDA(cudaMemcpyAsync(buf + N, b, lb * 32, cudaMemcpyHostToDevice, R.stream));
    if (int e = ntt_device<F>(buf, scratch, log_n, 2, false, nullptr)) return e;          // both forward transforms as a batch
    ntt_pointwise_mul_kernel<F><<<(unsigned)((N + 255) / 256), 256, 0, R.stream>>>(scratch, scratch + N, N);
    R.launches += 1;
    if (int e = ntt_device<F>(scratch, buf, log_n, 1, true, nullptr)) return e;
    HALO_CUDA(cudaMemcpyAsync(out, buf, lo * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
}
template int poly_mul_host<0>(const uint64_t*, size_t, const uint64_t*, size_t, uint64_t*);
template int poly_mul_host<1>(const uint64_t*, size_t, const uint64_t*, size_t, uint64_t*);

}  // namespace halo
