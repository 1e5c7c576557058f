// Variable-base multi-scalar multiplication on sm_100a: signed-digit Pippenger.
//
// Replaces ark-ec's `VariableBaseMSM::msm_unchecked` as reached from
//   crates/group/src/group.rs:48-56 (point_dot_affine / point_dot),
//   crates/accumulation/src/pedersen.rs:21 (commit), crates/accumulation/src/pcdl.rs:413,417 (IPA L/R).
// The result is a group element, so it is bit-identical (as canonical affine coordinates) to the reference's
// for any correct schedule.  Pipeline for k points, window width c, nwin = floor(255/c)+1 windows,
// B = 2^(c-1) buckets per window (signed digits in [-2^(c-1), 2^(c-1)]):
//   1. msm_count      scalars -> canonical -> signed digits; histogram of (window,bucket) sizes     [HBM: 32 B/pt]
//   2. msm_scan       exclusive prefix sum of the nwin*B bucket sizes
//   3. msm_scatter    digits again; write point index|sign into its bucket's slice                  [4*nwin B/pt]
//   4. msm_size_*     counting sort of buckets by size so that a warp works on equally long buckets
//   5. msm_accumulate one thread per bucket: XYZZ accumulator += (+/-)base  (8M+2S each)            [the hot loop]
//   6. msm_reduce_a/b per window sum_b (b+1)*bucket_b: per-thread running sums over 32-bucket segments, then a
//                     block-wide suffix scan (no scalar multiplications, only additions and a few doublings)
//   7. host           Horner over the nwin window sums (255 doublings; a single GPU thread would take longer
//                     than the PCIe round trip)
#include "runtime.cuh"

namespace halo {

// ---------------------------------------------------------------------------------------------------------
// helpers
// ---------------------------------------------------------------------------------------------------------
template <int F> __device__ __forceinline__ Fe<F> load_fe(const Fe<F>* p) {
    Fe<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = __ldg(q), b = __ldg(q + 1);
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int F> __device__ __forceinline__ void store_fe(Fe<F>* p, const Fe<F>& v) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    q[1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}
template <int C> __device__ __forceinline__ Affine<C> load_affine(const Affine<C>* p) {
    Affine<C> r;
    r.x = load_fe(&p->x); r.y = load_fe(&p->y);
    return r;
}
template <int C> __device__ __forceinline__ XYZZ<C> load_xyzz(const XYZZ<C>* p) {
    XYZZ<C> r;
    r.X = load_fe(&p->X); r.Y = load_fe(&p->Y); r.ZZ = load_fe(&p->ZZ); r.ZZZ = load_fe(&p->ZZZ);
    return r;
}
template <int C> __device__ __forceinline__ void store_xyzz(XYZZ<C>* p, const XYZZ<C>& v) {
    store_fe(&p->X, v.X); store_fe(&p->Y, v.Y); store_fe(&p->ZZ, v.ZZ); store_fe(&p->ZZZ, v.ZZZ);
}

// signed digit of window w (width c) of a canonical scalar, given the carry from the window below
__device__ __forceinline__ int window_digit(const u32 (&s)[8], int w, int c, int& carry) {
    int bit = w * c;
    int limb = bit >> 5, off = bit & 31;
    u64 lo = limb < 8 ? s[limb] : 0u;
    u64 hi = limb + 1 < 8 ? s[limb + 1] : 0u;
    u32 v = (u32)(((hi << 32) | lo) >> off) & ((1u << c) - 1u);
    int d = (int)v + carry;
    if (d > (1 << (c - 1))) { d -= (1 << c); carry = 1; } else carry = 0;
    return d;
}

// ---------------------------------------------------------------------------------------------------------
// 1. histogram of bucket sizes
// ---------------------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(256) msm_count(const Fe<SF>* __restrict__ scalars, size_t k, int c, int nwin,
                                                 u32* __restrict__ counts) {
    const u32 B = 1u << (c - 1);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < k; i += (size_t)gridDim.x * blockDim.x) {
        Fe<SF> s = fe_to_canonical<SF>(load_fe(scalars + i));
        int carry = 0;
        for (int w = 0; w < nwin; ++w) {
            int d = window_digit(s.l, w, c, carry);
            if (d != 0) atomicAdd(&counts[(size_t)w * B + (u32)(abs(d) - 1)], 1u);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 2. exclusive scan (single block; n up to a few million 32-bit counters)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) msm_scan(const u32* __restrict__ in, u32* __restrict__ out, size_t n) {
    __shared__ u32 part[1024];
    const int t = threadIdx.x;
    const size_t chunk = (n + 1023) / 1024;
    const size_t lo = min(n, (size_t)t * chunk), hi = min(n, lo + chunk);
    u32 s = 0;
    for (size_t i = lo; i < hi; ++i) s += in[i];
    part[t] = s;
    __syncthreads();
    // Hillis-Steele inclusive scan over the 1024 partials
    for (int d = 1; d < 1024; d <<= 1) {
        u32 v = t >= d ? part[t - d] : 0u;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    u32 run = t ? part[t - 1] : 0u;
    for (size_t i = lo; i < hi; ++i) { u32 v = in[i]; out[i] = run; run += v; }
}

// ---------------------------------------------------------------------------------------------------------
// 3. scatter point indices into bucket slices
// ---------------------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(256) msm_scatter(const Fe<SF>* __restrict__ scalars, size_t k, int c, int nwin,
                                                   const u32* __restrict__ offsets, u32* __restrict__ cursor,
                                                   u32* __restrict__ idx) {
    const u32 B = 1u << (c - 1);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < k; i += (size_t)gridDim.x * blockDim.x) {
        Fe<SF> s = fe_to_canonical<SF>(load_fe(scalars + i));
        int carry = 0;
        for (int w = 0; w < nwin; ++w) {
            int d = window_digit(s.l, w, c, carry);
            if (d != 0) {
                size_t b = (size_t)w * B + (u32)(abs(d) - 1);
                u32 pos = atomicAdd(&cursor[b], 1u);
                idx[offsets[b] + pos] = (u32)i | (d < 0 ? 0x80000000u : 0u);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 4. order buckets by size (descending): counting sort on min(size, 1023)
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) msm_size_hist(const u32* __restrict__ counts, size_t nb, u32* __restrict__ hist) {
    __shared__ u32 sh[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (size_t b = blockIdx.x * (size_t)blockDim.x + threadIdx.x; b < nb; b += (size_t)gridDim.x * blockDim.x)
        atomicAdd(&sh[1023u - min(counts[b], 1023u)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) if (sh[i]) atomicAdd(&hist[i], sh[i]);
}
__global__ void __launch_bounds__(256) msm_size_scatter(const u32* __restrict__ counts, size_t nb,
                                                        const u32* __restrict__ hist_off, u32* __restrict__ hist_cur,
                                                        u32* __restrict__ order) {
    for (size_t b = blockIdx.x * (size_t)blockDim.x + threadIdx.x; b < nb; b += (size_t)gridDim.x * blockDim.x) {
        u32 bin = 1023u - min(counts[b], 1023u);
        u32 pos = atomicAdd(&hist_cur[bin], 1u);
        order[hist_off[bin] + pos] = (u32)b;
    }
}

// ---------------------------------------------------------------------------------------------------------
// 5. bucket accumulation: one thread per bucket, buckets visited in size order
// ---------------------------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(128, 4) msm_accumulate(const Affine<C>* __restrict__ bases,
                                                         const u32* __restrict__ idx,
                                                         const u32* __restrict__ offsets,
                                                         const u32* __restrict__ counts,
                                                         const u32* __restrict__ order, size_t nb,
                                                         XYZZ<C>* __restrict__ buckets) {
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= nb) return;
    const u32 b = order[t];
    const u32 cnt = counts[b];
    const u32* my = idx + offsets[b];
    XYZZ<C> acc = xyzz_identity<C>();
    if (cnt) {
        u32 e = __ldg(my);
        Affine<C> p = load_affine(bases + (e & 0x7fffffffu));
        for (u32 j = 0; j < cnt; ++j) {
            // prefetch the next base while this addition runs
            u32 e_next = 0; Affine<C> p_next;
            if (j + 1 < cnt) { e_next = __ldg(my + j + 1); p_next = load_affine(bases + (e_next & 0x7fffffffu)); }
            xyzz_madd<C>(acc, p, (e >> 31) != 0);
            e = e_next; p = p_next;
        }
    }
    store_xyzz(buckets + b, acc);
}

// ---------------------------------------------------------------------------------------------------------
// 6a. per-segment running sums: segment of S consecutive buckets -> (sum, sum_j j*bucket_{lo+j})
// ---------------------------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(128, 2) msm_reduce_a(const XYZZ<C>* __restrict__ buckets, size_t nseg_total, int S,
                                                       XYZZ<C>* __restrict__ segsum, XYZZ<C>* __restrict__ segwgt) {
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= nseg_total) return;
    const XYZZ<C>* base = buckets + t * (size_t)S;
    XYZZ<C> run = xyzz_identity<C>(), acc = xyzz_identity<C>();
    for (int j = S - 1; j >= 1; --j) {
        XYZZ<C> bkt = load_xyzz(base + j);
        xyzz_add<C>(run, bkt);
        xyzz_add<C>(acc, run);
    }
    XYZZ<C> b0 = load_xyzz(base);
    xyzz_add<C>(run, b0);
    store_xyzz(segsum + t, run);
    store_xyzz(segwgt + t, acc);
}

// ---------------------------------------------------------------------------------------------------------
// 6b. one block per window: items j < nitems with (sum_j, wgt_j), item j covering buckets [j*S, (j+1)*S).
//     window result = sum_j [ wgt_j + (j*S) * sum_j ] + sum_j sum_j          (bucket b carries weight b+1)
// ---------------------------------------------------------------------------------------------------------
template <int C> __device__ __forceinline__ void xyzz_dbl_n(XYZZ<C>& p, int n) {
    for (int i = 0; i < n; ++i) p = xyzz_dbl<C>(p);
}

template <int C>
__global__ void __launch_bounds__(256, 1) msm_reduce_b(const XYZZ<C>* __restrict__ sums, const XYZZ<C>* __restrict__ wgts,
                                                       int nitems, int log2S, XYZZ<C>* __restrict__ winsum) {
    __shared__ XYZZ<C> sh[256];
    const int t = threadIdx.x, T = blockDim.x;
    const int w = blockIdx.x;
    sums += (size_t)w * nitems;
    if (wgts) wgts += (size_t)w * nitems;
    const int I = (nitems + T - 1) / T;                 // items per thread (power of two or 1)
    int log2I = 0; while ((1 << log2I) < I) ++log2I;
    const int lo = t * I;
    XYZZ<C> run = xyzz_identity<C>(), acc = xyzz_identity<C>(), wsum = xyzz_identity<C>();
    for (int j = I - 1; j >= 0; --j) {
        int it = lo + j;
        if (it < nitems) {
            XYZZ<C> s = load_xyzz(sums + it);
            xyzz_add<C>(run, s);
            if (wgts) { XYZZ<C> wg = load_xyzz(wgts + it); xyzz_add<C>(wsum, wg); }
        }
        if (j >= 1) xyzz_add<C>(acc, run);
    }
    // thread-local weighted sum relative to the thread's first bucket: wsum + S * acc
    xyzz_dbl_n<C>(acc, log2S);
    xyzz_add<C>(wsum, acc);
    // suffix scan of the thread sums
    sh[t] = run;
    __syncthreads();
    for (int d = 1; d < T; d <<= 1) {
        XYZZ<C> mine = sh[t];
        if (t + d < T) { XYZZ<C> o = sh[t + d]; xyzz_add<C>(mine, o); }
        __syncthreads();
        sh[t] = mine;
        __syncthreads();
    }
    // sum_t t * s_t = sum_{t>=1} suffix_t ;  thread t starts at bucket t*I*S
    XYZZ<C> z = wsum;
    if (t >= 1) {
        XYZZ<C> y = sh[t];
        xyzz_dbl_n<C>(y, log2S + log2I);
        xyzz_add<C>(z, y);
    } else {
        XYZZ<C> tot = sh[0];          // + sum of all buckets (weight b+1 instead of b)
        xyzz_add<C>(z, tot);
    }
    __syncthreads();
    sh[t] = z;
    __syncthreads();
    for (int d = T >> 1; d >= 1; d >>= 1) {
        if (t < d) { XYZZ<C> a = sh[t], b = sh[t + d]; xyzz_add<C>(a, b); sh[t] = a; }
        __syncthreads();
    }
    if (t == 0) store_xyzz(winsum + w, sh[0]);
}

// ---------------------------------------------------------------------------------------------------------
// host orchestration
// ---------------------------------------------------------------------------------------------------------
static int pick_window(size_t k) {
    int ov = rt().msm_window_override;
    if (ov >= 2 && ov <= 16) return ov;
    int lg = 0; while (((size_t)1 << (lg + 1)) <= k) ++lg;
    int c = lg - 4;
    if (c < 4) c = 4;
    if (c > 16) c = 16;
    return c;
}

struct MsmPlan {
    int c, nwin, S, log2S, nitems;
    size_t B, nb, nseg_total;
};
static MsmPlan make_plan(size_t k) {
    MsmPlan p;
    p.c = pick_window(k);
    p.nwin = 255 / p.c + 1;
    p.B = (size_t)1 << (p.c - 1);
    p.nb = p.B * p.nwin;
    p.S = p.B > 1024 ? (int)(p.B / 1024) : 1;
    p.log2S = 0; while ((1 << p.log2S) < p.S) ++p.log2S;
    p.nitems = (int)(p.B / p.S);
    p.nseg_total = (size_t)p.nitems * p.nwin;
    return p;
}

// Enqueues the whole device pipeline on the runtime stream and leaves nwin window sums (XYZZ) in *d_winsum.
template <int C>
static int msm_enqueue(const Affine<C>* d_bases, const Fe<CurveTraits<C>::SCALAR>* d_scalars, size_t k,
                       const MsmPlan& P, XYZZ<C>** d_winsum_out) {
    constexpr int SF = CurveTraits<C>::SCALAR;
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    u32 *counts, *offsets, *cursor, *idx, *order, *sizehist;
    XYZZ<C> *buckets, *segsum, *segwgt, *winsum;
    // counts | cursor | sizehist(3*1024) share one zero-filled allocation
    size_t zero_words = 2 * P.nb + 3 * 1024;
    if (int e = ws_get(WS_COUNTS, zero_words * 4, (void**)&counts)) return e;
    cursor = counts + P.nb;
    sizehist = cursor + P.nb;                       // [0,1024) hist, [1024,2048) offsets, [2048,3072) cursors
    if (int e = ws_get(WS_OFFSETS, P.nb * 4, (void**)&offsets)) return e;
    if (int e = ws_get(WS_IDX, (size_t)P.nwin * k * 4, (void**)&idx)) return e;
    if (int e = ws_get(WS_ORDER, P.nb * 4, (void**)&order)) return e;
    if (int e = ws_get(WS_BUCKETS, P.nb * sizeof(XYZZ<C>), (void**)&buckets)) return e;
    if (int e = ws_get(WS_SEGSUM, P.nseg_total * sizeof(XYZZ<C>), (void**)&segsum)) return e;
    if (int e = ws_get(WS_SEGWGT, P.nseg_total * sizeof(XYZZ<C>), (void**)&segwgt)) return e;
    if (int e = ws_get(WS_WINSUM, (size_t)P.nwin * sizeof(XYZZ<C>), (void**)&winsum)) return e;
    if (int e = prof_begin()) return e;
    HALO_CUDA(cudaMemsetAsync(counts, 0, zero_words * 4, st));

    int grid_pts = (int)std::min<size_t>((k + 255) / 256, (size_t)R.sm_count * 8);
    int grid_b = (int)std::min<size_t>((P.nb + 255) / 256, (size_t)R.sm_count * 8);
    msm_count<SF><<<grid_pts, 256, 0, st>>>(d_scalars, k, P.c, P.nwin, counts);
    prof_mark("count");
    msm_scan<<<1, 1024, 0, st>>>(counts, offsets, P.nb);
    prof_mark("scan");
    msm_scatter<SF><<<grid_pts, 256, 0, st>>>(d_scalars, k, P.c, P.nwin, offsets, cursor, idx);
    prof_mark("scatter");
    msm_size_hist<<<grid_b, 256, 0, st>>>(counts, P.nb, sizehist);
    msm_scan<<<1, 1024, 0, st>>>(sizehist, sizehist + 1024, 1024);
    msm_size_scatter<<<grid_b, 256, 0, st>>>(counts, P.nb, sizehist + 1024, sizehist + 2048, order);
    prof_mark("order");
    msm_accumulate<C><<<(unsigned)((P.nb + 127) / 128), 128, 0, st>>>(d_bases, idx, offsets, counts, order, P.nb, buckets);
    prof_mark("accumulate");
    R.launches += 8;
    if (P.S > 1) {
        msm_reduce_a<C><<<(unsigned)((P.nseg_total + 127) / 128), 128, 0, st>>>(buckets, P.nseg_total, P.S, segsum, segwgt);
        msm_reduce_b<C><<<P.nwin, 256, 0, st>>>(segsum, segwgt, P.nitems, P.log2S, winsum);
        R.launches += 1;
    } else {
        msm_reduce_b<C><<<P.nwin, 256, 0, st>>>(buckets, nullptr, P.nitems, 0, winsum);
    }
    prof_mark("reduce");
    HALO_CUDA(cudaGetLastError());
    *d_winsum_out = winsum;
    return HALO_OK;
}

// Horner over the window sums on the host: out = sum_w 2^(c*w) * win[w]
template <int C>
static void msm_combine_host(const XYZZ<C>* win, const MsmPlan& P, uint64_t out_xyz[12]) {
    XYZZ<C> acc = xyzz_identity<C>();
    for (int w = P.nwin - 1; w >= 0; --w) {
        for (int i = 0; i < P.c; ++i) acc = xyzz_dbl<C>(acc);
        xyzz_add<C>(acc, win[w]);
    }
    Jac<C> j = xyzz_to_jac<C>(acc);
    memcpy(out_xyz, &j, 96);
}

template <int C>
int msm_device(const void* d_bases, const void* d_scalars, size_t k, uint64_t out_xyz[12]) {
    Runtime& R = rt();
    if (k == 0) {
        Jac<C> j = xyzz_to_jac<C>(xyzz_identity<C>());
        memcpy(out_xyz, &j, 96);
        return HALO_OK;
    }
    HALO_CHECK(k < ((size_t)1 << 31), HALO_EINVAL, "msm: k must be < 2^31");
    MsmPlan P = make_plan(k);
    XYZZ<C>* d_win = nullptr;
    if (int e = msm_enqueue<C>((const Affine<C>*)d_bases, (const Fe<CurveTraits<C>::SCALAR>*)d_scalars, k, P, &d_win)) return e;
    void* h = nullptr;
    if (int e = pinned_get((size_t)P.nwin * sizeof(XYZZ<C>), &h)) return e;
    HALO_CUDA(cudaMemcpyAsync(h, d_win, (size_t)P.nwin * sizeof(XYZZ<C>), cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    msm_combine_host<C>((const XYZZ<C>*)h, P, out_xyz);
    return HALO_OK;
}
template int msm_device<0>(const void*, const void*, size_t, uint64_t*);
template int msm_device<1>(const void*, const void*, size_t, uint64_t*);

}  // namespace halo
