// Variable-base multi-scalar multiplication on sm_100a: signed-digit Pippenger.
//
// Replaces ark-ec's `VariableBaseMSM::msm_unchecked` as reached from
//   crates/group/src/group.rs:48-56 (point_dot_affine / point_dot),
//   crates/accumulation/src/pedersen.rs:21 (commit), crates/accumulation/src/pcdl.rs:413,417 (IPA L/R),
// and, batched over one resident base set, the 16 commits of a prover round
//   (crates/plonk/src/plonk/protocol.rs:114,263).
// The result is a group element, so it is bit-identical (as canonical affine coordinates) to the reference's
// for any correct schedule.  Pipeline for nbatch scalar vectors of k points over the same bases, window width c,
// nwin = floor(255/c)+1 windows per MSM, B = 2^(c-1) buckets per window (signed digits in [-2^(c-1), 2^(c-1)]);
// "window" below means one of the nbatch*nwin (MSM, window) pairs:
//   1. msm_count       scalars -> canonical -> signed digits; histogram of (window,bucket) sizes    [HBM: 32 B/pt]
//   2. scan_*          exclusive prefix sum of the bucket sizes (three small multi-block kernels)
//   3. msm_scatter     digits again; write point index|sign into its bucket's slice                 [4*nwin B/pt]
//   4. msm_task_*      cut every bucket into tasks of at most L points and counting-sort the tasks by size, so
//                      that a warp works on equally long tasks and no thread ever owns more than L points
//                      (degenerate scalar distributions put a whole window into a handful of buckets)
//   5. msm_accumulate  one thread per task: XYZZ accumulator += (+/-)base  (8M+2S each)             [the hot loop]
//   6. msm_fixup       buckets that were cut into several tasks: one warp sums the partial results
//   7. msm_reduce_level  per window sum_b (b+1)*bucket_b by recursive grouping (groups of 8 buckets carry a
//                      plain sum and a weighted sum up one level; only additions and a few doublings)
//   8. host            Horner over the nwin window sums of each MSM (255 doublings; a single GPU thread would
//                      take longer than the 2 KB D2H copy)
// MSMs of at most 2^13 points skip all of this for ONE launch (msm_small_kernel, further down) unless their
// scalars are badly skewed.
#include "runtime.cuh"
#include <algorithm>
#include <atomic>
#include <functional>
#include <thread>

namespace halo {

// ipa.cu: the commit-path table of the resident generators (rows 2^(c w) * Gs[i]); *tab == nullptr when there is none
template <int C> int ensure_fixed_table(const void** tab, int* c, int* ncols, size_t* row_stride);

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
// 1. histogram of bucket sizes.  scalars: nbatch vectors of k elements; counts: nbatch*nwin*B
// ---------------------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(256) msm_count(const Fe<SF>* __restrict__ scalars, size_t k, size_t total, int c,
                                                 int nwin, int fixed, u32* __restrict__ counts) {
    // fixed != 0 (fixed-base table path): the nwin digit columns of an MSM share ONE bucket window
    const u32 B = 1u << (c - 1);
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t j = t / k;                                   // which MSM of the batch
        Fe<SF> s = fe_to_canonical<SF>(load_fe(scalars + t));
        u32* cw = counts + j * (size_t)(fixed ? 1 : nwin) * B;
        int carry = 0;
        for (int w = 0; w < nwin; ++w) {
            int d = window_digit(s.l, w, c, carry);
            if (d != 0) atomicAdd(&cw[(size_t)(fixed ? 0 : w) * B + (u32)(abs(d) - 1)], 1u);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 2. exclusive scan of n 32-bit counters: per-block sums, scan of the block sums, apply.  1024 items / block.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ u32 block_exclusive_scan_256(u32 v, u32* sh /*[8]*/, u32& block_total) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    u32 x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { u32 y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
    if (lane == 31) sh[wid] = x;
    __syncthreads();
    u32 woff = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { u32 s = sh[i]; if (i < wid) woff += s; tot += s; }
    block_total = tot;
    __syncthreads();
    return woff + x - v;
}
__global__ void __launch_bounds__(256) scan_block_sums(const u32* __restrict__ in, size_t n, u32* __restrict__ bsum) {
    __shared__ u32 sh[8];
    size_t base = (size_t)blockIdx.x * 1024 + threadIdx.x * 4;
    u32 v = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) if (base + i < n) v += in[base + i];
    u32 tot;
    block_exclusive_scan_256(v, sh, tot);
    if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}
// single block: exclusive scan of up to 256*per items in place, also writes the grand total to *total_out
__global__ void __launch_bounds__(256) scan_partials(u32* __restrict__ bsum, size_t nblk, u32* __restrict__ total_out) {
    __shared__ u32 sh[8];
    const size_t per = (nblk + 255) / 256;
    const size_t lo = min(nblk, threadIdx.x * per), hi = min(nblk, lo + per);
    u32 v = 0;
    for (size_t i = lo; i < hi; ++i) v += bsum[i];
    u32 tot;
    u32 run = block_exclusive_scan_256(v, sh, tot);
    for (size_t i = lo; i < hi; ++i) { u32 x = bsum[i]; bsum[i] = run; run += x; }
    if (threadIdx.x == 0 && total_out) *total_out = tot;
}
// out = exclusive scan; out2 (optional) gets a second copy (used as the scatter cursor)
__global__ void __launch_bounds__(256) scan_apply(const u32* __restrict__ in, size_t n, const u32* __restrict__ bsum,
                                                  u32* __restrict__ out, u32* __restrict__ out2) {
    __shared__ u32 sh[8];
    size_t base = (size_t)blockIdx.x * 1024 + threadIdx.x * 4;
    u32 x[4], v = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { x[i] = base + i < n ? in[base + i] : 0u; v += x[i]; }
    u32 tot;
    u32 run = block_exclusive_scan_256(v, sh, tot) + bsum[blockIdx.x];
#pragma unroll
    for (int i = 0; i < 4; ++i) if (base + i < n) { out[base + i] = run; if (out2) out2[base + i] = run; run += x[i]; }
}

// ---------------------------------------------------------------------------------------------------------
// 3. scatter point indices into bucket slices.  cursor starts as a copy of the offsets.
// ---------------------------------------------------------------------------------------------------------
template <int SF>
__global__ void __launch_bounds__(256) msm_scatter(const Fe<SF>* __restrict__ scalars, size_t k, size_t total, int c,
                                                   int nwin, int fixed, u32 row_stride, u32* __restrict__ cursor, u32* __restrict__ idx) {
    // fixed != 0: one bucket window per MSM; the entry of digit column w is the table index w * row_stride + i
    // (row w of the table holds 2^(c w) * G_i)
    const u32 B = 1u << (c - 1);
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const size_t j = t / k;
        const u32 i = (u32)(t - j * k);
        Fe<SF> s = fe_to_canonical<SF>(load_fe(scalars + t));
        u32* cw = cursor + j * (size_t)(fixed ? 1 : nwin) * B;
        int carry = 0;
        for (int w = 0; w < nwin; ++w) {
            int d = window_digit(s.l, w, c, carry);
            if (d != 0) {
                u32 pos = atomicAdd(&cw[(size_t)(fixed ? 0 : w) * B + (u32)(abs(d) - 1)], 1u);
                idx[pos] = (fixed ? (u32)w * row_stride + i : i) | (d < 0 ? 0x80000000u : 0u);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 4. tasks: bucket b with cnt points becomes max(1, ceil(cnt/L)) tasks of at most L points; tasks are
//    counting-sorted by size (descending: bin = L - size).  A bucket cut into several tasks ("multi") reserves a
//    run of slots in the partial-result array.
// ---------------------------------------------------------------------------------------------------------
struct TaskMeta { u32 ntasks, npartial, nmulti, nchunks, nbig, pad[3]; };
constexpr u32 FIXUP_CHUNK = 256;          // partial results one warp adds in the first fix-up pass
constexpr int MSM_MAX_L = 512;

__global__ void __launch_bounds__(256) msm_task_hist(const u32* __restrict__ counts, size_t nb, u32 L,
                                                     u32* __restrict__ hist, TaskMeta* __restrict__ meta,
                                                     u32* __restrict__ pslot, u32* __restrict__ cbase,
                                                     uint2* __restrict__ chunks, u32* __restrict__ big) {
    __shared__ u32 sh[MSM_MAX_L + 1];
    for (int i = threadIdx.x; i <= MSM_MAX_L; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (size_t b = blockIdx.x * (size_t)blockDim.x + threadIdx.x; b < nb; b += (size_t)gridDim.x * blockDim.x) {
        const u32 cnt = counts[b];
        const u32 full = cnt / L, rem = cnt - full * L;
        if (full) atomicAdd(&sh[0], full);
        if (rem || cnt == 0) atomicAdd(&sh[L - rem], 1u);
        const u32 nt = full + (rem ? 1u : 0u);
        if (nt > 1) {
            pslot[b] = atomicAdd(&meta->npartial, nt);
            // fix-up work list: one warp per chunk of FIXUP_CHUNK partial results
            const u32 nch = (nt + FIXUP_CHUNK - 1) / FIXUP_CHUNK;
            const u32 c0 = atomicAdd(&meta->nchunks, nch);
            cbase[b] = c0;
            for (u32 q = 0; q < nch; ++q) chunks[c0 + q] = make_uint2((u32)b, q);
            if (nch > 1) big[atomicAdd(&meta->nbig, 1u)] = (u32)b;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i <= (int)L; i += blockDim.x) if (sh[i]) atomicAdd(&hist[i], sh[i]);
}
// bincur: exclusive scan of hist (in place), used as cursors.  The L + 1 cursors are hit by every bucket, so the
// atomics are aggregated per warp: lanes with the same bin elect a leader (match.any), full-length tasks go through a
// warp prefix sum -- one atomic per (warp, bin) instead of one per bucket.
__global__ void __launch_bounds__(256) msm_task_emit(const u32* __restrict__ counts, size_t nb, u32 L,
                                                     u32* __restrict__ bincur, uint2* __restrict__ tasks) {
    const u32 lane = threadIdx.x & 31;
    for (size_t base = blockIdx.x * (size_t)blockDim.x; base < nb; base += (size_t)gridDim.x * blockDim.x) {
        const size_t b = base + threadIdx.x;
        const bool valid = b < nb;
        const u32 cnt = valid ? counts[b] : 0u;
        const u32 full = cnt / L, rem = cnt - full * L;
        // full-length tasks: bin 0
        u32 incl = full;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { u32 y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= (u32)d) incl += y; }
        const u32 warp_full = __shfl_sync(0xffffffffu, incl, 31);
        u32 pos0 = 0;
        if (warp_full) {
            if (lane == 31) pos0 = atomicAdd(&bincur[0], warp_full);
            pos0 = __shfl_sync(0xffffffffu, pos0, 31) + incl - full;
            for (u32 s = 0; s < full; ++s) tasks[pos0 + s] = make_uint2((u32)b, s);
        }
        // the remainder task (or the single empty task of an empty bucket): bin L - rem
        const bool has = valid && (rem || cnt == 0);
        const u32 key = has ? L - rem : 0xffffffffu;
        const u32 peers = __match_any_sync(0xffffffffu, key);
        const u32 leader = (u32)__ffs((int)peers) - 1u;
        u32 pos = 0;
        if (has && lane == leader) pos = atomicAdd(&bincur[key], (u32)__popc(peers));
        pos = __shfl_sync(0xffffffffu, pos, leader);
        if (has) tasks[pos + (u32)__popc(peers & ((1u << lane) - 1u))] = make_uint2((u32)b, full);
    }
}

// ---------------------------------------------------------------------------------------------------------
// 5. bucket accumulation: one thread per task, tasks visited in size order
// ---------------------------------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(128, 4) msm_accumulate(const Affine<C>* __restrict__ bases,
                                                         const u32* __restrict__ idx,
                                                         const u32* __restrict__ offsets,
                                                         const u32* __restrict__ counts,
                                                         const uint2* __restrict__ tasks,
                                                         const TaskMeta* __restrict__ meta, u32 L,
                                                         const u32* __restrict__ pslot,
                                                         XYZZ<C>* __restrict__ buckets, XYZZ<C>* __restrict__ partial,
                                                         u32 buckets_per_msm, ptrdiff_t base_stride, int into) {
    // into != 0: the bucket array already holds the sums of earlier point ranges of the same MSM (msm_host_chunked):
    // a bucket's task starts from the stored value instead of the identity
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= meta->ntasks) return;
    const uint2 tk = tasks[t];
    const u32 b = tk.x, seg = tk.y;
    const u32 cnt_b = counts[b];
    const u32 cnt = min(L, cnt_b - seg * L);
    const u32* my = idx + offsets[b] + seg * L;
    XYZZ<C>* dest = cnt_b > L ? partial + pslot[b] + seg : buckets + b;
    if (base_stride) bases += (ptrdiff_t)(b / buckets_per_msm) * base_stride;      // MSM j of a batch reads bases + j*stride
    XYZZ<C> acc = (into && cnt_b <= L) ? load_xyzz(buckets + b) : xyzz_identity<C>();
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
    store_xyzz(dest, acc);
}

// ---------------------------------------------------------------------------------------------------------
// 5b. bucket accumulation with BATCHED AFFINE additions (large problems).  An affine addition costs 1 inversion +
//     2M + 1S; Montgomery's trick shares one inversion among the independent additions of a batch at 3M each, so an
//     addition is 5M + 1S = 6 multiplications instead of the 8M + 2S of the XYZZ mixed addition -- provided the
//     inversion is cheap and the batch large.  The inversion is fe_inv_safegcd (~1/5 of a Fermat ladder); the batch
//     is built per THREAD: a thread owns G consecutive positions of the bucket-sorted index array (whatever buckets
//     they belong to) and reduces every bucket segment in it pairwise, round after round: round 0 adds positions
//     (0,1), (2,3), ... of each segment, round 1 the results, and so on -- G/2, G/4, ... independent additions per
//     round, one inversion per round.  When a round would have fewer than BA_MIN_PAIRS additions the inversion no
//     longer pays and the remaining few items per segment are added with XYZZ mixed additions.  Intermediate points
//     live in a thread-private, lane-interleaved scratch (a warp's accesses to slot q are one contiguous run).
//     A segment that is a whole bucket goes to the bucket array; a bucket that continues across a thread boundary
//     leaves "head" / "tail" partial sums, added up by msm_ba_fixup.  Every case of the group law is handled
//     (identity inputs, P + P, P - P): the reference's generator set is duplicate-heavy.
//     No thread owns more than G points whatever the scalar distribution, so the task list is not needed.
//     MEASURED (profiles/r2g_msm_accumulate_ba_ncu.txt): correct on every edge case, but slower than the task path on
//     this field -- 3.17 ms vs 2.56 ms at 2^20 points.  32 % fewer wide multiplies, yet only 13 % fewer instructions
//     (inversions, planning, the second pass over the operands), 3.3 warps per scheduler, and every addition waits
//     for operands it has just requested (long-scoreboard 2.5 stalls per issue) where the task path prefetches its
//     next base.  Kept as an opt-in path (tuning key 8); the default is the task path.
// ---------------------------------------------------------------------------------------------------------
constexpr int BA_T = 128;                 // threads per CTA
constexpr u32 BA_MAX_G = 256;             // positions per thread (masks are 256 bits)
constexpr u32 BA_MIN_PAIRS = 24;          // below this an inversion (~50 multiplications) costs more than it saves

__device__ __forceinline__ u32 upper_bound_u32(const u32* __restrict__ a, u32 n, u32 v) {   // first i with a[i] > v
    u32 lo = 0, hi = n;
    while (lo < hi) { const u32 mid = (lo + hi) >> 1; if (__ldg(a + mid) <= v) lo = mid + 1; else hi = mid; }
    return lo;
}
template <int F> __device__ __forceinline__ Fe<F> ld_fe_plain(const Fe<F>* p) {       // scratch written by this kernel: no __ldg
    Fe<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = q[0], b = q[1];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w; r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
// denominator of P1 + P2 and the case: 0 = chord (x2 - x1), 1 = tangent (2 y1), 2 = no inversion needed (an identity
// input, or P + (-P))
template <int C>
__device__ __forceinline__ int ba_pair_case(const Affine<C>& p1, const Affine<C>& p2, Fe<CurveTraits<C>::BASE>& den) {
    constexpr int B = CurveTraits<C>::BASE;
    if (affine_is_identity<C>(p1) || affine_is_identity<C>(p2)) { den = fe_one<B>(); return 2; }
    if (fe_eq<B>(p1.x, p2.x)) {
        if (fe_eq<B>(p1.y, p2.y)) { den = fe_dbl<B>(p1.y); return 1; }     // y != 0: the curves have no point of order 2
        den = fe_one<B>();
        return 2;
    }
    den = fe_sub<B>(p2.x, p1.x);
    return 0;
}

template <int C>
__global__ void __launch_bounds__(BA_T, 4) msm_accumulate_ba(const Affine<C>* __restrict__ bases, const u32* __restrict__ idx,
                                                             const u32* __restrict__ offsets, const u32* __restrict__ counts,
                                                             u32 nb, u32 G, Affine<C>* Wbase, Fe<CurveTraits<C>::BASE>* Pbase,
                                                             XYZZ<C>* __restrict__ buckets, XYZZ<C>* __restrict__ head,
                                                             XYZZ<C>* __restrict__ tail) {
    constexpr int B = CurveTraits<C>::BASE;
    __shared__ u32 s_mask[8][BA_T], s_new[8][BA_T], s_pair[8][BA_T];     // per-thread 256-bit masks, conflict-free columns
    const u32 tid = threadIdx.x, lane = tid & 31u;
    const u32 total = __ldg(offsets + nb - 1) + __ldg(counts + nb - 1);
    const u32 per_chunk = BA_T * G;
    const u32 nchunks = (u32)(((u64)total + per_chunk - 1) / per_chunk);
    const size_t warp_id = (size_t)blockIdx.x * (BA_T / 32) + (tid >> 5);
    Affine<C>* W = Wbase + warp_id * (size_t)G * 32 + lane;               // slot q of this thread: W[q * 32]
    Fe<B>* P = Pbase + warp_id * (size_t)(G / 2) * 32 + lane;             // prefix product of pair j: P[j * 32]
#define BA_BIT(m, i) ((m[(i) >> 5][tid] >> ((i) & 31u)) & 1u)
#define BA_SET(m, i) (m[(i) >> 5][tid] |= 1u << ((i) & 31u))
    for (u32 chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
        const u32 vt = chunk * BA_T + tid;                                // virtual thread: positions [vt G, vt G + G)
        const u64 p0w = (u64)vt * G;
        if (p0w >= total) continue;
        const u32 p0 = (u32)p0w, p1 = (u32)(p0w + G < total ? p0w + G : total);
        store_xyzz(head + vt, xyzz_identity<C>());
        store_xyzz(tail + vt, xyzz_identity<C>());
        const u32 b0 = upper_bound_u32(offsets, nb, p0) - 1u;             // the (non-empty) bucket holding position p0
#pragma unroll
        for (int w = 0; w < 8; ++w) s_mask[w][tid] = 0u;
        s_mask[0][tid] = 1u;
        for (u32 b = b0 + 1; b < nb; ++b) {                               // segment starts inside (p0, p1)
            const u32 o = __ldg(offsets + b);
            if (o >= p1) break;
            BA_SET(s_mask, o - p0);
        }
        u32 n = p1 - p0, round = 0;
        auto load_pt = [&](u32 a) -> Affine<C> {
            Affine<C> pt;
            if (round == 0) {
                const u32 e = __ldg(idx + p0 + a);
                pt = load_affine(bases + (e & 0x7fffffffu));
                if (e >> 31) pt.y = fe_neg<B>(pt.y);                      // -(0,0) stays (0,0)
            } else {
                pt.x = ld_fe_plain(&W[(size_t)a * 32].x); pt.y = ld_fe_plain(&W[(size_t)a * 32].y);
            }
            return pt;
        };
        for (;;) {
            // plan the round: output slot q takes items (a, a+1) of one segment, or a lone last item
#pragma unroll
            for (int w = 0; w < 8; ++w) { s_new[w][tid] = 0u; s_pair[w][tid] = 0u; }
            u32 np = 0, q = 0, a = 0;
            while (a < n) {
                const bool pr = a + 1 < n && !BA_BIT(s_mask, a + 1);
                if (BA_BIT(s_mask, a)) BA_SET(s_new, q);
                if (pr) { BA_SET(s_pair, q); ++np; }
                a += 1u + (pr ? 1u : 0u);
                ++q;
            }
            const u32 nout = q;
            if (np < BA_MIN_PAIRS) break;
            // pass 1, slots in descending order: prefix products of the denominators
            Fe<B> run = fe_one<B>();
            u32 j = 0;
            a = n;
            for (q = nout; q-- > 0;) {
                const bool pr = BA_BIT(s_pair, q) != 0u;
                a -= 1u + (pr ? 1u : 0u);
                if (!pr) continue;
                Fe<B> den;
                Fe<B> x1, x2;
                if (round == 0) {
                    x1 = load_fe(&bases[__ldg(idx + p0 + a) & 0x7fffffffu].x);
                    x2 = load_fe(&bases[__ldg(idx + p0 + a + 1) & 0x7fffffffu].x);
                } else {
                    x1 = ld_fe_plain(&W[(size_t)a * 32].x); x2 = ld_fe_plain(&W[(size_t)(a + 1) * 32].x);
                }
                if (fe_eq<B>(x1, x2) || fe_is_zero<B>(x1) || fe_is_zero<B>(x2)) {          // rare: look at the whole points
                    const Affine<C> q1 = load_pt(a), q2 = load_pt(a + 1);
                    ba_pair_case<C>(q1, q2, den);
                } else den = fe_sub<B>(x2, x1);
                store_fe(&P[(size_t)j * 32], run);
                run = fe_mul<B>(run, den);
                ++j;
            }
            Fe<B> inv = fe_inv_safegcd<B>(run);
            // pass 2, ascending: 1/den_j = inv * prefix_j, then inv *= den_j; results overwrite the scratch in place
            // (slot q <= first input a of every slot, so nothing still needed is overwritten)
            a = 0;
            for (q = 0; q < nout; ++q) {
                const bool pr = BA_BIT(s_pair, q) != 0u;
                if (!pr) {
                    if (round == 0 || a != q) {
                        const Affine<C> pt = load_pt(a);
                        store_fe(&W[(size_t)q * 32].x, pt.x); store_fe(&W[(size_t)q * 32].y, pt.y);
                    }
                    a += 1;
                    continue;
                }
                --j;
                const Affine<C> q1 = load_pt(a), q2 = load_pt(a + 1);
                Fe<B> den;
                const int cs = ba_pair_case<C>(q1, q2, den);
                Affine<C> r;
                if (cs == 2) {
                    if (affine_is_identity<C>(q1)) r = q2;
                    else if (affine_is_identity<C>(q2)) r = q1;
                    else { r.x = fe_zero<B>(); r.y = fe_zero<B>(); }
                } else {
                    const Fe<B> dinv = fe_mul<B>(inv, ld_fe_plain(&P[(size_t)j * 32]));
                    inv = fe_mul<B>(inv, den);
                    Fe<B> num;
                    if (cs == 0) num = fe_sub<B>(q2.y, q1.y);
                    else { const Fe<B> xx = fe_sqr<B>(q1.x); num = fe_add<B>(fe_dbl<B>(xx), xx); }
                    const Fe<B> lam = fe_mul<B>(num, dinv);
                    r.x = fe_sub<B>(fe_sub<B>(fe_sqr<B>(lam), q1.x), q2.x);
                    r.y = fe_sub<B>(fe_mul<B>(lam, fe_sub<B>(q1.x, r.x)), q1.y);
                }
                store_fe(&W[(size_t)q * 32].x, r.x); store_fe(&W[(size_t)q * 32].y, r.y);
                a += 2;
            }
#pragma unroll
            for (int w = 0; w < 8; ++w) s_mask[w][tid] = s_new[w][tid];
            n = nout;
            ++round;
        }
        // the few items left per segment: XYZZ mixed additions, then out to the bucket / the boundary partial sums
        {
            XYZZ<C> acc = xyzz_identity<C>();
            u32 b = b0;
            bool first = true;
            auto flush = [&](bool last) {
                const u32 ob = __ldg(offsets + b);
                const bool is_head = first && ob < p0;
                const bool is_tail = last && (u64)ob + __ldg(counts + b) > (u64)p1;
                store_xyzz(is_head ? head + vt : is_tail ? tail + vt : buckets + b, acc);
            };
            for (u32 a = 0; a < n; ++a) {
                if (a > 0 && BA_BIT(s_mask, a)) {
                    flush(false);
                    first = false;
                    do { ++b; } while (__ldg(counts + b) == 0u);
                    acc = xyzz_identity<C>();
                }
                const Affine<C> pt = load_pt(a);
                xyzz_madd<C>(acc, pt, false);
            }
            flush(true);
        }
    }
#undef BA_BIT
#undef BA_SET
}

template <int C> __device__ __forceinline__ XYZZ<C> xyzz_shfl_down(const XYZZ<C>& v, int d);
template <int C> __device__ __forceinline__ XYZZ<C> warp_sum_xyzz(XYZZ<C> acc, u32 live);
// Thread boundaries v*G inside a bucket: the bucket is tail[v-1] + head[v] + ... + head[last].  Only the FIRST boundary
// inside a bucket acts.  The sum is almost always two terms: one THREAD per boundary adds runs of up to BA_FIX_SHORT
// terms; longer runs (skewed scalar distributions: a bucket spanning many threads) are queued for msm_ba_fixup_long,
// one warp per run.
constexpr u32 BA_FIX_SHORT = 8;
template <int C>
__global__ void __launch_bounds__(128) msm_ba_fixup(const u32* __restrict__ offsets, const u32* __restrict__ counts, u32 nb, u32 G,
                                                    const XYZZ<C>* __restrict__ head, const XYZZ<C>* __restrict__ tail,
                                                    XYZZ<C>* __restrict__ buckets, u32* __restrict__ nlong, u32* __restrict__ longs) {
    const u32 total = __ldg(offsets + nb - 1) + __ldg(counts + nb - 1);
    const u32 VT = (u32)(((u64)total + G - 1) / G);
    const u64 v = (u64)blockIdx.x * blockDim.x + threadIdx.x + 1;
    if (v >= VT) return;
    const u64 pos = v * G;
    const u32 b = upper_bound_u32(offsets, nb, (u32)(pos - 1)) - 1u;      // bucket holding the position before the boundary
    const u64 ob = __ldg(offsets + b), end = ob + __ldg(counts + b);
    if (end <= pos) return;                                               // the bucket ends at the boundary
    if (ob < (v - 1) * G) return;                                         // an earlier boundary lies inside it too
    const u64 last = (end - 1) / G;
    if (last - (v - 1) + 1 > BA_FIX_SHORT) { longs[atomicAdd(nlong, 1u)] = (u32)v; return; }
    XYZZ<C> acc = load_xyzz(tail + (v - 1));
    for (u64 u = v; u <= last; ++u) { const XYZZ<C> x = load_xyzz(head + u); xyzz_add<C>(acc, x); }
    store_xyzz(buckets + b, acc);
}
template <int C>
__global__ void __launch_bounds__(128) msm_ba_fixup_long(const u32* __restrict__ offsets, const u32* __restrict__ counts, u32 nb, u32 G,
                                                         const XYZZ<C>* __restrict__ head, const XYZZ<C>* __restrict__ tail,
                                                         XYZZ<C>* __restrict__ buckets, const u32* __restrict__ nlong,
                                                         const u32* __restrict__ longs) {
    const u32 lane = threadIdx.x & 31u;
    const u32 nwarps = gridDim.x * (blockDim.x >> 5), n = *nlong;
    for (u32 w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < n; w += nwarps) {
        const u64 v = longs[w], pos = v * G;
        const u32 b = upper_bound_u32(offsets, nb, (u32)(pos - 1)) - 1u;
        const u64 end = (u64)__ldg(offsets + b) + __ldg(counts + b), last = (end - 1) / G;
        XYZZ<C> acc = xyzz_identity<C>();
        for (u64 u = v - 1 + lane; u <= last; u += 32) {
            const XYZZ<C> x = load_xyzz(u == v - 1 ? tail + u : head + u);
            xyzz_add<C>(acc, x);
        }
        acc = warp_sum_xyzz<C>(acc, 32u);
        if (lane == 0) store_xyzz(buckets + b, acc);
    }
}

// ---------------------------------------------------------------------------------------------------------
// 6. multi-task buckets: one warp adds the partial results of one bucket
// ---------------------------------------------------------------------------------------------------------
template <int C> __device__ __forceinline__ XYZZ<C> xyzz_shfl_down(const XYZZ<C>& v, int d) {
    XYZZ<C> r;
    const u32* s = reinterpret_cast<const u32*>(&v);
    u32* o = reinterpret_cast<u32*>(&r);
#pragma unroll
    for (int i = 0; i < 32; ++i) o[i] = __shfl_down_sync(0xffffffffu, s[i], d);
    return r;
}
template <int C> __device__ __forceinline__ XYZZ<C> warp_sum_xyzz(XYZZ<C> acc, u32 live) {
    int d0 = 16;
    while (d0 >= 1 && (u32)d0 >= live) d0 >>= 1;            // lanes >= live hold the identity
    for (int d = d0; d >= 1; d >>= 1) { XYZZ<C> o = xyzz_shfl_down<C>(acc, d); xyzz_add<C>(acc, o); }
    return acc;
}
// pass 1: one warp per chunk of <= FIXUP_CHUNK partial results; a bucket with a single chunk is finished here
template <int C>
__global__ void __launch_bounds__(128, 2) msm_fixup(const uint2* __restrict__ chunks, const TaskMeta* __restrict__ meta,
                                                    const u32* __restrict__ counts, const u32* __restrict__ pslot,
                                                    const u32* __restrict__ cbase, u32 L, const XYZZ<C>* __restrict__ partial,
                                                    XYZZ<C>* __restrict__ partial2, XYZZ<C>* __restrict__ buckets, int into) {
    const u32 lane = threadIdx.x & 31;
    const u32 nwarps = gridDim.x * (blockDim.x >> 5);
    const u32 nchunks = meta->nchunks;
    for (u32 w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < nchunks; w += nwarps) {
        const uint2 ch = chunks[w];
        const u32 b = ch.x, q = ch.y;
        const u32 nt = (counts[b] + L - 1) / L;
        const u32 lo = q * FIXUP_CHUNK, hi = min(nt, lo + FIXUP_CHUNK);
        const XYZZ<C>* src = partial + pslot[b];
        XYZZ<C> acc = xyzz_identity<C>();
        for (u32 s = lo + lane; s < hi; s += 32) { XYZZ<C> v = load_xyzz(src + s); xyzz_add<C>(acc, v); }
        acc = warp_sum_xyzz<C>(acc, hi - lo);
        if (lane == 0) {
            if (into && nt <= FIXUP_CHUNK) { const XYZZ<C> old = load_xyzz(buckets + b); xyzz_add<C>(acc, old); }
            store_xyzz(nt <= FIXUP_CHUNK ? buckets + b : partial2 + cbase[b] + q, acc);
        }
    }
}
// pass 2: buckets with more than FIXUP_CHUNK partial results: one warp adds the chunk sums
template <int C>
__global__ void __launch_bounds__(128, 2) msm_fixup2(const u32* __restrict__ big, const TaskMeta* __restrict__ meta,
                                                     const u32* __restrict__ counts, const u32* __restrict__ cbase, u32 L,
                                                     const XYZZ<C>* __restrict__ partial2, XYZZ<C>* __restrict__ buckets, int into) {
    const u32 lane = threadIdx.x & 31;
    const u32 nwarps = gridDim.x * (blockDim.x >> 5);
    const u32 nbig = meta->nbig;
    for (u32 m = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); m < nbig; m += nwarps) {
        const u32 b = big[m];
        const u32 nt = (counts[b] + L - 1) / L;
        const u32 nch = (nt + FIXUP_CHUNK - 1) / FIXUP_CHUNK;
        const XYZZ<C>* src = partial2 + cbase[b];
        XYZZ<C> acc = xyzz_identity<C>();
        for (u32 s = lane; s < nch; s += 32) { XYZZ<C> v = load_xyzz(src + s); xyzz_add<C>(acc, v); }
        acc = warp_sum_xyzz<C>(acc, nch);
        if (lane == 0) {
            if (into) { const XYZZ<C> old = load_xyzz(buckets + b); xyzz_add<C>(acc, old); }
            store_xyzz(buckets + b, acc);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// 7. window sums  W = sum_b (b+1) * bucket_b  as a binary tree with pending plain sums.
//    Let s(l)_i be the sum of 2^l consecutive buckets.  Then  sum_b b*bucket_b = sum_l 2^l * A_l  with
//    A_l = sum over ODD i of s(l)_i, and  sum_b bucket_b = T = s(c-1)_0.  Level kernel l halves the main array
//    (s(l+1)_g = s(l)_2g + s(l)_2g+1), starts the pending array of A_l with the odd items, and halves every older
//    pending array: (l+1) * n/2^(l+1) independent additions per window and a dependency depth of ONE addition per
//    level (c-1 levels), instead of a chain of 2S-1 additions per level of a grouped running sum.  The last kernel
//    doubles A_l l times and adds everything up inside one warp per window.
// ---------------------------------------------------------------------------------------------------------
template <int C> __device__ __forceinline__ void xyzz_dbl_n(XYZZ<C>& p, int n) {
    for (int i = 0; i < n; ++i) p = xyzz_dbl<C>(p);
}
template <int C>
__global__ void __launch_bounds__(128, 3) msm_tree_level(const XYZZ<C>* __restrict__ main_in, const XYZZ<C>* __restrict__ pend_in,
                                                         XYZZ<C>* __restrict__ main_out, XYZZ<C>* __restrict__ pend_out,
                                                         u32 n_in, u32 lvl, size_t nthreads) {
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= nthreads) return;
    const u32 n_out = n_in >> 1;
    const size_t per_win = (size_t)(lvl + 1) * n_out;
    const size_t win = t / per_win;
    const u32 rem = (u32)(t - win * per_win);
    const u32 r = rem / n_out, g = rem - r * n_out;
    const XYZZ<C>* src = r < lvl ? pend_in + (win * lvl + r) * (size_t)n_in : main_in + win * (size_t)n_in;
    XYZZ<C> a = load_xyzz(src + 2 * g), b = load_xyzz(src + 2 * g + 1);
    XYZZ<C>* pout = pend_out + (win * (lvl + 1) + r) * (size_t)n_out + g;
    if (r == lvl) store_xyzz(pout, b);                  // odd item of this level: first summand of A_lvl
    xyzz_add<C>(a, b);
    if (r < lvl) store_xyzz(pout, a);
    else store_xyzz(main_out + win * (size_t)n_out + g, a);
}
// one warp per window: lane r < nlev holds A_r (weight 2^r), lane nlev holds T
template <int C>
__global__ void __launch_bounds__(128, 2) msm_tree_final(const XYZZ<C>* __restrict__ main_in, const XYZZ<C>* __restrict__ pend_in,
                                                         u32 nlev, size_t nwt, XYZZ<C>* __restrict__ winsum) {
    const u32 lane = threadIdx.x & 31;
    const size_t win = blockIdx.x * (size_t)(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (win >= nwt) return;
    XYZZ<C> v = xyzz_identity<C>();
    if (lane < nlev) { v = load_xyzz(pend_in + win * nlev + lane); xyzz_dbl_n<C>(v, (int)lane); }
    else if (lane == nlev) v = load_xyzz(main_in + win);
    for (int d = 16; d >= 1; d >>= 1) { XYZZ<C> o = xyzz_shfl_down<C>(v, d); xyzz_add<C>(v, o); }
    if (lane == 0) store_xyzz(winsum + win, v);
}

// ---------------------------------------------------------------------------------------------------------
// host orchestration
// ---------------------------------------------------------------------------------------------------------
// Window width.  Only widths whose TOP window is well populated are used: 255 = 15*16+15 = 17*15 = 19*13+8 =
// 31*8+7, whereas c = 14, 12, 11, 9 ... leave a 2-3 bit top window whose handful of buckets receive every point
// (same-address atomics in count/scatter, long fix-up chains).  Break-even sizes measured on B200 (tools/msm_tune.py).
static int pick_window(size_t k) {
    int ov = rt().msm_window_override;
    if (ov >= 2 && ov <= 16) return ov;
    int lg = 0; while (((size_t)1 << (lg + 1)) <= k) ++lg;
    if (lg >= 20) return 16;
    if (lg >= 17) return 15;
    if (lg >= 15) return 13;
    return 8;
}

struct MsmPlan {
    int c, nwin;          // window width, windows per MSM
    size_t nbatch, nwt;   // MSMs in the batch, total windows = nbatch*nwin
    size_t B, nb;         // buckets per window, total buckets
    u32 L;                // max points per task
    size_t max_tasks, max_partial;
    u32 ba_G;             // > 0: batched-affine accumulation with this many positions per thread (5b)
    // fixed-base table path (msm_fixed_device): ncols digit columns per MSM share one bucket window (nwin == 1), the
    // bases are rows of 2^(c w) * G_i of row_stride points each
    int ncols = 0;        // 0: the ordinary pipeline (digit columns == windows)
    u32 row_stride = 0;
};
static MsmPlan make_plan(size_t k, size_t nbatch) {
    MsmPlan p;
    p.c = pick_window(k);
    p.nwin = 255 / p.c + 1;
    p.nbatch = nbatch;
    p.nwt = nbatch * p.nwin;
    p.B = (size_t)1 << (p.c - 1);
    p.nb = p.B * p.nwt;
    // long tasks amortise the bucket write; short tasks give small problems enough threads
    size_t work = k * p.nwt;
    p.L = work >= ((size_t)1 << 23) ? 128 : work >= ((size_t)1 << 20) ? 32 : 16;
    if (work >= ((size_t)1 << 23)) {                     // keep ordinary buckets whole: L >= 4 x the mean bucket size
        size_t mean4 = 4 * (k / p.B);
        if (mean4 > p.L) p.L = (u32)std::min<size_t>(mean4, MSM_MAX_L);
    }
    p.max_tasks = p.nb + work / p.L + 1;
    p.max_partial = 2 * (work / p.L) + 2;
    // Batched-affine accumulation is OPT-IN (tuning key 8 = 64 / 128 / 256 positions per thread): measured on B200 at
    // 2^20 points it takes 3.17 ms against 2.56 ms for the XYZZ task path (only 13 % fewer
    // instructions per addition on this 255-bit field, and latency-bound where the task path is pipe-bound).
    const int t8 = rt().tune[8];
    p.ba_G = (t8 == 64 || t8 == 128 || t8 == 256) ? (u32)t8 : 0u;
    return p;
}

// plan of the fixed-base table path: ncols digit columns of width c per MSM into ONE window of 2^(c-1) buckets
static MsmPlan make_plan_fixed(size_t k, size_t nbatch, int c, int ncols, size_t row_stride) {
    MsmPlan p;
    p.c = c;
    p.nwin = 1;
    p.nbatch = nbatch;
    p.nwt = nbatch;
    p.B = (size_t)1 << (c - 1);
    p.nb = p.B * p.nwt;
    const size_t work = k * (size_t)ncols * nbatch;
    p.L = 128;
    const size_t mean4 = 4 * (k * (size_t)ncols / p.B);
    if (mean4 > p.L) p.L = (u32)std::min<size_t>(mean4, MSM_MAX_L);
    p.max_tasks = p.nb + work / p.L + 1;
    p.max_partial = 2 * (work / p.L) + 2;
    p.ba_G = 0;
    p.ncols = ncols;
    p.row_stride = (u32)row_stride;
    return p;
}

static int scan_u32(const u32* in, size_t n, u32* out, u32* out2, u32* bsum /*>= n/1024+1*/, u32* total_out, cudaStream_t st) {
    unsigned nblk = (unsigned)((n + 1023) / 1024);
    scan_block_sums<<<nblk, 256, 0, st>>>(in, n, bsum);
    scan_partials<<<1, 256, 0, st>>>(bsum, nblk, total_out);
    scan_apply<<<nblk, 256, 0, st>>>(in, n, bsum, out, out2);
    rt().launches += 3;
    return HALO_OK;
}

// Enqueues the whole device pipeline on the runtime stream and leaves nwt window sums (XYZZ) in *d_winsum_out.
template <int C>
static int msm_enqueue(const Affine<C>* d_bases, const Fe<CurveTraits<C>::SCALAR>* d_scalars, size_t k,
                       const MsmPlan& P, ptrdiff_t base_stride, XYZZ<C>** d_winsum_out, int into = 0, bool reduce = true) {
    // into: accumulate on top of the bucket array's contents; reduce = false: stop after the buckets (more point
    // ranges of the same MSM follow)
    constexpr int SF = CurveTraits<C>::SCALAR;
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    const size_t total = k * P.nbatch;
    u32 *counts, *offsets, *cursor, *bsum, *idx, *pslot, *cbase, *big, *hist;
    uint2 *tasks, *chunks;
    TaskMeta* meta;
    XYZZ<C> *buckets, *partial, *partial2, *lvl[4], *winsum;
    // counts | hist (MSM_MAX_L+1, padded to 1024) | meta share one zero-filled allocation
    const size_t zero_words = P.nb + 1024 + 8;
    if (int e = ws_get(WS_COUNTS, zero_words * 4, (void**)&counts)) return e;
    hist = counts + P.nb;
    meta = (TaskMeta*)(hist + 1024);
    const size_t nblk = (P.nb + 1023) / 1024;
    if (int e = ws_get(WS_OFFSETS, (2 * P.nb + nblk + 1) * 4, (void**)&offsets)) return e;
    cursor = offsets + P.nb;
    bsum = cursor + P.nb;
    const int dcols = P.ncols ? P.ncols : P.nwin;          // digit columns per MSM
    if (int e = ws_get(WS_IDX, P.nbatch * (size_t)dcols * k * 4, (void**)&idx)) return e;
    if (int e = ws_get(WS_ORDER, P.max_tasks * sizeof(uint2), (void**)&tasks)) return e;
    // pslot | cbase | big (nb words each) | chunk list (max_partial entries)
    if (int e = ws_get(WS_SIZEHIST, 3 * P.nb * 4 + P.max_partial * sizeof(uint2), (void**)&pslot)) return e;
    cbase = pslot + P.nb;
    big = cbase + P.nb;
    chunks = (uint2*)(big + P.nb);
    if (int e = ws_get(WS_BUCKETS, P.nb * sizeof(XYZZ<C>), (void**)&buckets)) return e;
    partial = partial2 = nullptr;
    if (!P.ba_G) {
        if (int e = ws_get(WS_TMP3, 2 * P.max_partial * sizeof(XYZZ<C>), (void**)&partial)) return e;
        partial2 = partial + P.max_partial;
    }
    // tree reduction: main ping-pong (nb/2, nb/4) and pending ping-pong (nb/2 each)
    const size_t half = P.nb / 2 + P.nwt;
    if (int e = ws_get(WS_SEGSUM, (half + half / 2 + 1) * sizeof(XYZZ<C>), (void**)&lvl[0])) return e;
    lvl[1] = lvl[0] + half;
    if (int e = ws_get(WS_SEGWGT, 2 * half * sizeof(XYZZ<C>), (void**)&lvl[2])) return e;
    lvl[3] = lvl[2] + half;
    if (int e = ws_get(WS_WINSUM, P.nwt * sizeof(XYZZ<C>), (void**)&winsum)) return e;

    if (!into) if (int e = prof_begin()) return e;
    HALO_CUDA(cudaMemsetAsync(counts, 0, zero_words * 4, st));
    const int grid_pts = (int)std::min<size_t>((total + 255) / 256, (size_t)R.sm_count * 8);
    const int grid_b = (int)std::min<size_t>((P.nb + 255) / 256, (size_t)R.sm_count * 8);
    msm_count<SF><<<grid_pts, 256, 0, st>>>(d_scalars, k, total, P.c, dcols, P.ncols ? 1 : 0, counts);
    prof_mark("count");
    scan_u32(counts, P.nb, offsets, cursor, bsum, nullptr, st);
    prof_mark("scan");
    msm_scatter<SF><<<grid_pts, 256, 0, st>>>(d_scalars, k, total, P.c, dcols, P.ncols ? 1 : 0, P.row_stride, cursor, idx);
    prof_mark("scatter");
    if (P.ba_G) {
        // batched-affine accumulation (5b): persistent grid, scratch per resident warp
        const unsigned grid = (unsigned)R.sm_count * 4;
        const size_t warps = (size_t)grid * (BA_T / 32);
        Affine<C>* W; Fe<CurveTraits<C>::BASE>* Pp; XYZZ<C>* head;
        if (int e = ws_get(WS_BA_W, warps * P.ba_G * 32 * sizeof(Affine<C>), (void**)&W)) return e;
        if (int e = ws_get(WS_BA_P, warps * (P.ba_G / 2) * 32 * 32, (void**)&Pp)) return e;
        const size_t vt_max = (P.nwt * k + P.ba_G - 1) / P.ba_G + 1;
        if (int e = ws_get(WS_TMP3, 2 * vt_max * sizeof(XYZZ<C>), (void**)&head)) return e;
        HALO_CUDA(cudaMemsetAsync(buckets, 0, P.nb * sizeof(XYZZ<C>), st));          // empty buckets are the identity
        prof_mark("tasks");
        if (int e = msm_wait_bases(st)) return e;                                     // bases uploaded on copy_stream
        msm_accumulate_ba<C><<<grid, BA_T, 0, st>>>(d_bases, idx, offsets, counts, (u32)P.nb, P.ba_G, W, Pp, buckets, head, head + vt_max);
        prof_mark("accumulate");
        // the queue of long runs: counter in the zero-filled meta block, entries in the (otherwise unused) `big` array
        msm_ba_fixup<C><<<(unsigned)((vt_max + 127) / 128), 128, 0, st>>>(offsets, counts, (u32)P.nb, P.ba_G, head, head + vt_max, buckets,
                                                                        &meta->nbig, big);
        msm_ba_fixup_long<C><<<R.sm_count, 128, 0, st>>>(offsets, counts, (u32)P.nb, P.ba_G, head, head + vt_max, buckets, &meta->nbig, big);
        prof_mark("fixup");
        R.launches += 6;
    } else {
    msm_task_hist<<<grid_b, 256, 0, st>>>(counts, P.nb, P.L, hist, meta, pslot, cbase, chunks, big);
    scan_partials<<<1, 256, 0, st>>>(hist, P.L + 1, &meta->ntasks);
    msm_task_emit<<<grid_b, 256, 0, st>>>(counts, P.nb, P.L, hist, tasks);
    prof_mark("tasks");
    if (int e = msm_wait_bases(st)) return e;                                                         // bases uploaded on copy_stream
    msm_accumulate<C><<<(unsigned)((P.max_tasks + 127) / 128), 128, 0, st>>>(d_bases, idx, offsets, counts, tasks, meta,
                                                                           P.L, pslot, buckets, partial,
                                                                           (u32)(P.B * P.nwin), base_stride, into);
    prof_mark("accumulate");
    msm_fixup<C><<<R.sm_count * 4, 128, 0, st>>>(chunks, meta, counts, pslot, cbase, P.L, partial, partial2, buckets, into);
    msm_fixup2<C><<<R.sm_count, 128, 0, st>>>(big, meta, counts, cbase, P.L, partial2, buckets, into);
    prof_mark("fixup");
    R.launches += 8;
    }
    if (!reduce) { HALO_CUDA(cudaGetLastError()); return HALO_OK; }
    // bucket reduction: c-1 tree levels, then the per-window combine
    {
        const XYZZ<C>*m_in = buckets, *p_in = nullptr;
        u32 n_in = (u32)P.B, lv = 0;
        while (n_in > 1) {
            XYZZ<C>* m_out = lvl[lv & 1];
            XYZZ<C>* p_out = lvl[2 + (lv & 1)];
            const size_t threads = (size_t)(lv + 1) * (n_in >> 1) * P.nwt;
            msm_tree_level<C><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(m_in, p_in, m_out, p_out, n_in, lv, threads);
            R.launches += 1;
            m_in = m_out; p_in = p_out; n_in >>= 1; ++lv;
        }
        msm_tree_final<C><<<(unsigned)((P.nwt + 3) / 4), 128, 0, st>>>(m_in, p_in, lv, P.nwt, winsum);
        R.launches += 1;
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

// ---------------------------------------------------------------------------------------------------------
// Small MSMs (k <= 2^13) in ONE launch.  The general pipeline above is ~25 dependent launches, which is a
// 0.45-0.8 ms latency floor; the IPA spends its last 15 rounds on MSM pairs of this size (crates/accumulation/src/
// pcdl.rs:413,417 once the vectors are short) and small proofs consist of nothing else.  Here one CTA owns one
// (MSM, 8-bit window, point range) triple end to end: signed digits -> counting sort of the point indices by bucket in
// shared memory -> four threads per bucket accumulate its run -> (b+1)*bucket_b by double-and-add -> tree sum.
// The host adds the range partials and runs the same Horner as for the general path.  A window whose largest
// bucket is far above the mean (degenerate scalar distributions) raises a flag and the caller falls back to the
// task-balanced general pipeline.
// ---------------------------------------------------------------------------------------------------------
constexpr int SMALL_C = 8, SMALL_B = 128, SMALL_NWIN = 32, SMALL_T = 512;
constexpr size_t SMALL_MAX_K = 8192;      // measured crossover against the general pipeline: between 2^13 and 2^14

// signed digit (width 8) of window w of a canonical scalar s < 2^255: with A = 0x7f in every byte, the carry into
// window w under the rule "digit > 128 -> digit - 256, carry 1" is the carry into bit 8w of s + A
__device__ __forceinline__ int small_digit(const u32 (&s)[8], u32 w) {
    const u32 limb = w >> 2, sh = (w & 3u) * 8u;
    u32 carry = 0, t = 0;
#pragma unroll
    for (u32 i = 0; i < 8; ++i) {
        u64 x = (u64)s[i] + 0x7f7f7f7fu + carry;
        if (i == limb) t = (u32)x;
        carry = (u32)(x >> 32);
    }
    u32 sl = 0;
#pragma unroll
    for (u32 i = 0; i < 8; ++i) if (i == limb) sl = s[i];
    const u32 cin = ((t ^ sl ^ 0x7f7f7f7fu) >> sh) & 1u;
    int d = (int)((sl >> sh) & 0xffu) + (int)cin;
    if (d > 128) d -= 256;
    return d;
}

template <int C>
__global__ void __launch_bounds__(SMALL_T, 1) msm_small_kernel(const Affine<C>* __restrict__ bases,
                                                               const Fe<CurveTraits<C>::SCALAR>* __restrict__ scalars, u32 k,
                                                               ptrdiff_t base_stride, u32 split, u32 chunk, u32 max_bucket,
                                                               XYZZ<C>* __restrict__ out, u32* __restrict__ flag) {
    constexpr int SF = CurveTraits<C>::SCALAR;
    extern __shared__ __align__(16) unsigned char small_raw[];
    unsigned short* sorted = reinterpret_cast<unsigned short*>(small_raw);   // <= 32768 entries: index in range | sign << 15
    XYZZ<C>* red = reinterpret_cast<XYZZ<C>*>(small_raw);                    // reused after the accumulation
    __shared__ u32 cnt[SMALL_B], off[SMALL_B], cur[SMALL_B], wsum[4], s_bad;
    const u32 tid = threadIdx.x;
    const u32 h = blockIdx.x % split, w = (blockIdx.x / split) % SMALL_NWIN, j = blockIdx.x / (split * SMALL_NWIN);
    const u32 lo = h * chunk, hi = min(k, lo + chunk);
    const Fe<SF>* sc = scalars + (size_t)j * k;
    bases += (ptrdiff_t)j * base_stride;
    if (tid < SMALL_B) cnt[tid] = 0;
    if (tid == 0) s_bad = 0;
    __syncthreads();
    for (u32 p = lo + tid; p < hi; p += SMALL_T) {
        Fe<SF> s = fe_to_canonical<SF>(load_fe(sc + p));
        int d = small_digit(s.l, w);
        if (d != 0) atomicAdd(&cnt[abs(d) - 1], 1u);
    }
    __syncthreads();
    if (tid < SMALL_B) {                                   // exclusive scan of the 128 bucket sizes (4 warps)
        const u32 v = cnt[tid];
        u32 x = v;
#pragma unroll
        for (int dlt = 1; dlt < 32; dlt <<= 1) { u32 y = __shfl_up_sync(0xffffffffu, x, dlt); if ((tid & 31) >= (u32)dlt) x += y; }
        if ((tid & 31) == 31) wsum[tid >> 5] = x;
        if (v > max_bucket) s_bad = 1;
        off[tid] = x - v;
    }
    __syncthreads();
    if (s_bad) { if (tid == 0) atomicOr(flag, 1u); return; }
    if (tid < SMALL_B) {
        u32 add = 0;
        for (u32 i = 0; i < (tid >> 5); ++i) add += wsum[i];
        off[tid] += add; cur[tid] = off[tid];
    }
    __syncthreads();
    for (u32 p = lo + tid; p < hi; p += SMALL_T) {
        Fe<SF> s = fe_to_canonical<SF>(load_fe(sc + p));
        int d = small_digit(s.l, w);
        if (d != 0) {
            u32 pos = atomicAdd(&cur[abs(d) - 1], 1u);
            sorted[pos] = (unsigned short)((p - lo) | (d < 0 ? 0x8000u : 0u));
        }
    }
    __syncthreads();
    // four threads per bucket, each a contiguous quarter of the bucket's run
    const u32 b = tid >> 2, q = tid & 3u;
    XYZZ<C> acc = xyzz_identity<C>();
    {
        const u32 n_b = cnt[b], o_b = off[b];
        const u32 e0 = o_b + (n_b * q) / 4, e1 = o_b + (n_b * (q + 1)) / 4;
        if (e0 < e1) {
            u32 e = sorted[e0];
            Affine<C> pt = load_affine(bases + lo + (e & 0x7fffu));
            for (u32 i = e0; i < e1; ++i) {
                u32 e_next = 0; Affine<C> p_next;
                if (i + 1 < e1) { e_next = sorted[i + 1]; p_next = load_affine(bases + lo + (e_next & 0x7fffu)); }
                xyzz_madd<C>(acc, pt, (e >> 15) != 0);
                e = e_next; pt = p_next;
            }
        }
    }
    { XYZZ<C> o = xyzz_shfl_down<C>(acc, 2); xyzz_add<C>(acc, o); }
    { XYZZ<C> o = xyzz_shfl_down<C>(acc, 1); xyzz_add<C>(acc, o); }
    __syncthreads();                                       // every thread is done with `sorted`
    // sum_b (b+1) * bucket_b = sum_i S_i with S_i = sum_{b >= i} bucket_b: a suffix scan (7 steps) and a tree sum (7)
    if (q == 0) red[b] = acc;
    __syncthreads();
    if (tid < SMALL_B) acc = red[tid];
    for (u32 st = 1; st < SMALL_B; st <<= 1) {
        const bool on = tid < SMALL_B && tid + st < SMALL_B;
        if (on) xyzz_add<C>(acc, red[tid + st]);
        __syncthreads();
        if (on) red[tid] = acc;
        __syncthreads();
    }
    for (u32 st = SMALL_B / 2; st >= 1; st >>= 1) {
        if (tid < st) { xyzz_add<C>(acc, red[tid + st]); red[tid] = acc; }
        __syncthreads();
    }
    if (tid == 0) store_xyzz(out + blockIdx.x, acc);
}

// returns HALO_OK with *done = false when the caller must use the general pipeline instead
template <int C>
static int msm_small(const Affine<C>* d_bases, const Fe<CurveTraits<C>::SCALAR>* d_scalars, size_t k, size_t nbatch,
                     ptrdiff_t base_stride, uint64_t* out_xyz, bool* done) {
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    *done = false;
    u32 split = (u32)std::max<size_t>(1, std::min<size_t>(4, (size_t)R.sm_count / (SMALL_NWIN * nbatch)));
    while (split > 1 && (k + split - 1) / split < 2048) --split;
    const u32 chunk = (u32)((k + split - 1) / split);
    const size_t ncta = nbatch * SMALL_NWIN * split;
    // a uniform window holds chunk/128 points per bucket (fewer when scalars are zero); tolerate 4x that
    const u32 max_bucket = 4 * (chunk / SMALL_B) + 64;
    static bool attr_set[2] = {false, false};
    const size_t smem = std::max<size_t>((size_t)chunk * 2, (size_t)SMALL_B * sizeof(XYZZ<C>));
    if (!attr_set[C]) {
        HALO_CUDA(cudaFuncSetAttribute(msm_small_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
        attr_set[C] = true;
    }
    unsigned char* buf;
    if (int e = ws_get(WS_WINSUM, 256 + ncta * sizeof(XYZZ<C>), (void**)&buf)) return e;
    u32* d_flag = (u32*)buf;
    XYZZ<C>* d_out = (XYZZ<C>*)(buf + 256);
    if (int e = prof_begin()) return e;
    if (int e = msm_wait_bases(st)) return e;                                                         // bases uploaded on copy_stream
    HALO_CUDA(cudaMemsetAsync(d_flag, 0, 4, st));
    msm_small_kernel<C><<<(unsigned)ncta, SMALL_T, smem, st>>>(d_bases, d_scalars, (u32)k, base_stride, split, chunk, max_bucket,
                                                              d_out, d_flag);
    R.launches += 1;
    prof_mark("small");
    HALO_CUDA(cudaGetLastError());
    void* hbuf = nullptr;
    if (int e = pinned_get(256 + ncta * sizeof(XYZZ<C>), &hbuf)) return e;
    HALO_CUDA(cudaMemcpyAsync(hbuf, buf, 256 + ncta * sizeof(XYZZ<C>), cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    if (*(const u32*)hbuf != 0) return HALO_OK;            // skewed window: general pipeline
    const XYZZ<C>* part = (const XYZZ<C>*)((const unsigned char*)hbuf + 256);
    MsmPlan P;
    P.c = SMALL_C; P.nwin = SMALL_NWIN;
    auto combine = [&](size_t i) {
        XYZZ<C> win[SMALL_NWIN];
        for (int w = 0; w < SMALL_NWIN; ++w) {
            const XYZZ<C>* src = part + (i * SMALL_NWIN + w) * split;
            win[w] = src[0];
            for (u32 hh = 1; hh < split; ++hh) xyzz_add<C>(win[w], src[hh]);
        }
        msm_combine_host<C>(win, P, out_xyz + 12 * i);
    };
    if (nbatch == 1) combine(0);
    else {
        Workers wk;
        wk.spawn_indexed(1, (unsigned)nbatch, [&](unsigned i) { combine(i); });
        combine(0);
        wk.join();
    }
    *done = true;
    return HALO_OK;
}

// ---------------------------------------------------------------------------------------------------------
// One MSM over HOST buffers, pipelined by point range.  halo_msm used to upload all scalars, sort, and then wait for
// ALL base points before the first addition: 96 B per point over the link sat in front of the accumulation.  Here the
// points are cut into NCH ranges; range i is uploaded (scalars, then bases) while range i-1 is being sorted and
// accumulated INTO THE SAME bucket array (`into` mode of the kernels above), and the bucket reduction runs once at
// the end.  Same window width and buckets as the unchunked MSM of k points, so the result is the same group element.
// Pinned or registered buffers: every copy asynchronous.  Pageable buffers (a Rust Vec): staged through pinned chunks
// by a helper thread (upload_staged) while this thread enqueues the kernels.  Four ranges in both cases.
// ---------------------------------------------------------------------------------------------------------
// The device half: leaves the window sums of the whole MSM in device memory (*d_winsum; *c_out / *nwin_out say how to
// combine them) with everything queued on the runtime stream -- the multi-GPU exchange (dist.cu) all-gathers them from
// there; msm_host_chunked below copies them back and combines them.
template <int C>
int msm_host_chunked_windows(const uint64_t* h_bases, const uint64_t* h_scalars, size_t k, void** d_winsum, int* c_out, int* nwin_out) {
    Runtime& R = rt();
    cudaStream_t st = R.stream;
    const bool pageable = host_ptr_is_pageable(h_bases) || host_ptr_is_pageable(h_scalars);
    const int t10 = R.tune[10];                                           // experiment knob: number of ranges (2..16)
    const int NCH = (t10 >= 2 && t10 <= 16) ? t10 : 4;                    // tools/e2e_chunks.py: flat from 2 to 8 (pinned), best at 4 (pageable)
    char *d_s, *d_b;
    if (int e = ws_get(WS_SCALARS, k * 32, (void**)&d_s)) return e;
    if (int e = ws_get(WS_BASES, k * 64, (void**)&d_b)) return e;
    for (int i = 0; i < 2 * NCH; ++i)
        if (!R.chunk_ev[i]) HALO_CUDA(cudaEventCreateWithFlags(&R.chunk_ev[i], cudaEventDisableTiming));
    HALO_CUDA(cudaEventRecord(R.copy_go, st));                            // the copies start behind what is already queued
    HALO_CUDA(cudaStreamWaitEvent(R.copy_stream, R.copy_go, 0));
    const MsmPlan P0 = make_plan(k, 1);
    HALO_CHECK(P0.nwt * k < ((size_t)1 << 32), HALO_EINVAL, "msm: too large (nwin*k must be < 2^32)");
    auto range = [&](int i, size_t& lo, size_t& hi) { lo = k * i / NCH; hi = k * (i + 1) / NCH; };
    std::atomic<int> recorded{0};                                         // events recorded so far (pageable path)
    std::atomic<int> up_rc{HALO_OK};
    Workers up;
    if (!pageable) {
        for (int i = 0; i < NCH; ++i) {
            size_t lo, hi; range(i, lo, hi);
            HALO_CUDA(cudaMemcpyAsync(d_s + lo * 32, h_scalars + lo * 4, (hi - lo) * 32, cudaMemcpyHostToDevice, R.copy_stream));
            HALO_CUDA(cudaEventRecord(R.chunk_ev[2 * i], R.copy_stream));
            HALO_CUDA(cudaMemcpyAsync(d_b + lo * 64, h_bases + lo * 8, (hi - lo) * 64, cudaMemcpyHostToDevice, R.copy_stream));
            HALO_CUDA(cudaEventRecord(R.chunk_ev[2 * i + 1], R.copy_stream));
        }
        recorded.store(2 * NCH);
    } else {
        up.spawn(1, [&]() {
            int rc = cudaSetDevice(R.device) == cudaSuccess ? HALO_OK : HALO_ECUDA;
            for (int i = 0; i < NCH && !rc; ++i) {
                size_t lo, hi; range(i, lo, hi);
                rc = upload_staged(d_s + lo * 32, h_scalars + lo * 4, (hi - lo) * 32, R.copy_stream);
                if (!rc && cudaEventRecord(R.chunk_ev[2 * i], R.copy_stream) != cudaSuccess) rc = HALO_ECUDA;
                if (!rc) recorded.store(2 * i + 1, std::memory_order_release);
                if (!rc) rc = upload_staged(d_b + lo * 64, h_bases + lo * 8, (hi - lo) * 64, R.copy_stream);
                if (!rc && cudaEventRecord(R.chunk_ev[2 * i + 1], R.copy_stream) != cudaSuccess) rc = HALO_ECUDA;
                if (!rc) recorded.store(2 * i + 2, std::memory_order_release);
            }
            if (rc) up_rc.store(rc);
        });
        HALO_CHECK(up.size() == 1, HALO_ECUDA, "msm: could not start the upload thread");
    }
    // wait (on the host) until event e has been recorded by the upload thread, then make `s` wait for it on the device
    auto wait_event = [&](int e, cudaStream_t s) -> int {
        while (recorded.load(std::memory_order_acquire) <= e) {
            if (up_rc.load()) { last_err() = "msm: staged upload failed"; return up_rc.load(); }
            std::this_thread::yield();
        }
        HALO_CUDA(cudaStreamWaitEvent(s, R.chunk_ev[e], 0));
        return HALO_OK;
    };
    XYZZ<C>* d_win = nullptr;
    int rc = HALO_OK;
    for (int i = 0; i < NCH && !rc; ++i) {
        size_t lo, hi; range(i, lo, hi);
        MsmPlan P = P0;                                                   // same windows and buckets, task sizes for this range
        const size_t kc = hi - lo, work = kc * P.nwt;
        P.L = work >= ((size_t)1 << 23) ? 128 : work >= ((size_t)1 << 20) ? 32 : 16;
        if (work >= ((size_t)1 << 23)) { const size_t mean4 = 4 * (kc / P.B); if (mean4 > P.L) P.L = (u32)std::min<size_t>(mean4, MSM_MAX_L); }
        P.max_tasks = P.nb + work / P.L + 1;
        P.max_partial = 2 * (work / P.L) + 2;
        P.ba_G = 0;
        rc = wait_event(2 * i, st);
        if (rc) break;
        R.bases_hook = [&, i](cudaStream_t s) { return wait_event(2 * i + 1, s); };
        rc = msm_enqueue<C>((const Affine<C>*)(d_b + lo * 64), (const Fe<CurveTraits<C>::SCALAR>*)(d_s + lo * 32), kc, P, 0, &d_win,
                            i > 0, i == NCH - 1);
        R.bases_hook = nullptr;
    }
    if (rc) up_rc.store(rc);                                              // lets the upload thread's waiters stop
    up.join();
    if (rc) { cudaStreamSynchronize(R.copy_stream); cudaStreamSynchronize(st); return rc; }
    *d_winsum = d_win; *c_out = P0.c; *nwin_out = P0.nwin;
    return HALO_OK;
}
template int msm_host_chunked_windows<0>(const uint64_t*, const uint64_t*, size_t, void**, int*, int*);
template int msm_host_chunked_windows<1>(const uint64_t*, const uint64_t*, size_t, void**, int*, int*);

template <int C>
int msm_host_chunked(const uint64_t* h_bases, const uint64_t* h_scalars, size_t k, uint64_t* out_xyz) {
    Runtime& R = rt();
    void* d_win = nullptr;
    MsmPlan P0;
    if (int e = msm_host_chunked_windows<C>(h_bases, h_scalars, k, &d_win, &P0.c, &P0.nwin)) return e;
    const size_t nwt = (size_t)P0.nwin;
    void* h = nullptr;
    if (int e = pinned_get(nwt * sizeof(XYZZ<C>), &h)) return e;
    HALO_CUDA(cudaMemcpyAsync(h, d_win, nwt * sizeof(XYZZ<C>), cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    msm_combine_host<C>((const XYZZ<C>*)h, P0, out_xyz);
    return HALO_OK;
}
template int msm_host_chunked<0>(const uint64_t*, const uint64_t*, size_t, uint64_t*);
template int msm_host_chunked<1>(const uint64_t*, const uint64_t*, size_t, uint64_t*);

// nbatch MSMs of k points; MSM j uses bases + j*base_stride (0: all share one base set); scalars: nbatch x k
// elements, out: nbatch x 12 words
template <int C>
int msm_device(const void* d_bases, const void* d_scalars, size_t k, size_t nbatch, uint64_t* out_xyz, ptrdiff_t base_stride) {
    Runtime& R = rt();
    if (k == 0 || nbatch == 0) {
        Jac<C> j = xyzz_to_jac<C>(xyzz_identity<C>());
        for (size_t i = 0; i < nbatch; ++i) memcpy(out_xyz + 12 * i, &j, 96);
        return HALO_OK;
    }
    HALO_CHECK(k < ((size_t)1 << 31), HALO_EINVAL, "msm: k must be < 2^31");
    if (k <= SMALL_MAX_K && nbatch <= 4 && R.tune[1] != 1) {       // one-launch path (tuning key 1 = 1 disables it)
        bool done = false;
        if (int e = msm_small<C>((const Affine<C>*)d_bases, (const Fe<CurveTraits<C>::SCALAR>*)d_scalars, k, nbatch, base_stride,
                                 out_xyz, &done)) return e;
        if (done) return HALO_OK;
    }
    MsmPlan P = make_plan(k, nbatch);
    if (base_stride) P.ba_G = 0;                          // per-MSM base sets (the IPA's L/R pair): task path
    // Commitments over the resident generators (pedersen::commit, pedersen.rs:7-27: the bases are pp.Gs[0..k)): the
    // fixed-base table of the parameter set turns the digit columns into ONE bucket window per MSM with c = 18-20
    // (tuning key 14 = 1: off).  Large prefixes only -- below half the set the ordinary pipeline is as good.
    // The L / R pair of an opening's FIRST round has the same property (its generators are still pp.Gs): two MSMs over
    // the two halves, addressed like the ordinary pair (bases + k with a stride of -k) inside every row of the table.
    const bool commit_shape = d_bases == R.pp[C].d_gs && base_stride == 0 && 2 * k >= R.pp[C].n && k <= R.pp[C].n;
    const bool lr_shape = nbatch == 2 && 2 * k == R.pp[C].n && base_stride == -(ptrdiff_t)k &&
                          d_bases == (const void*)((const Affine<C>*)R.pp[C].d_gs + k);
    if ((commit_shape || lr_shape) && R.tune[14] != 1 && R.tune[8] == 0 && R.pp[C].n >= ((size_t)1 << 18) && nbatch <= 64) {
        const void* tab = nullptr; int fc = 0, fcols = 0; size_t stride = 0;
        if (int e = ensure_fixed_table<C>(&tab, &fc, &fcols, &stride)) return e;
        if (tab && nbatch * (size_t)fcols * k < ((size_t)1 << 32) && (size_t)fcols * stride < ((size_t)1 << 31)) {
            P = make_plan_fixed(k, nbatch, fc, fcols, stride);
            d_bases = lr_shape ? (const void*)((const Affine<C>*)tab + k) : tab;
        }
    }
    HALO_CHECK(P.nwt * k < ((size_t)1 << 32), HALO_EINVAL, "msm: batch too large (nbatch*nwin*k must be < 2^32)");
    XYZZ<C>* d_win = nullptr;
    if (int e = msm_enqueue<C>((const Affine<C>*)d_bases, (const Fe<CurveTraits<C>::SCALAR>*)d_scalars, k, P, base_stride, &d_win)) return e;
    void* h = nullptr;
    if (int e = pinned_get(P.nwt * sizeof(XYZZ<C>), &h)) return e;
    HALO_CUDA(cudaMemcpyAsync(h, d_win, P.nwt * sizeof(XYZZ<C>), cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    const XYZZ<C>* win = (const XYZZ<C>*)h;
    if (nbatch == 1) {
        msm_combine_host<C>(win, P, out_xyz);
    } else {                                              // one Horner (255 doublings, ~55 us) per MSM: the calling thread takes a share
        unsigned nt = std::min<size_t>(nbatch, std::max(1u, std::thread::hardware_concurrency()));
        auto share = [&](unsigned t) { for (size_t i = t; i < nbatch; i += nt) msm_combine_host<C>(win + i * P.nwin, P, out_xyz + 12 * i); };
        Workers wk;
        wk.spawn_indexed(1, nt, share);
        share(0);
        wk.join();
    }
    return HALO_OK;
}
// The device half only, for the multi-GPU exchange (dist.cu): runs the general pipeline and leaves the nbatch * nwin
// window sums in device memory; *c_out / *nwin_out tell the receiver how to combine them.  k > 0.
template <int C>
int msm_windows_device(const void* d_bases, const void* d_scalars, size_t k, size_t nbatch, ptrdiff_t base_stride,
                       void** d_winsum, int* c_out, int* nwin_out) {
    HALO_CHECK(k > 0 && k < ((size_t)1 << 31), HALO_EINVAL, "msm: k must be in [1, 2^31)");
    MsmPlan P = make_plan(k, nbatch);
    if (base_stride) P.ba_G = 0;
    HALO_CHECK(P.nwt * k < ((size_t)1 << 32), HALO_EINVAL, "msm: batch too large (nbatch*nwin*k must be < 2^32)");
    XYZZ<C>* d_win = nullptr;
    if (int e = msm_enqueue<C>((const Affine<C>*)d_bases, (const Fe<CurveTraits<C>::SCALAR>*)d_scalars, k, P, base_stride, &d_win)) return e;
    *d_winsum = d_win; *c_out = P.c; *nwin_out = P.nwin;
    return HALO_OK;
}
template int msm_windows_device<0>(const void*, const void*, size_t, size_t, ptrdiff_t, void**, int*, int*);
template int msm_windows_device<1>(const void*, const void*, size_t, size_t, ptrdiff_t, void**, int*, int*);
// Horner over nwin window sums of width c (host)
template <int C>
void msm_horner_host(const void* win, int c, int nwin, uint64_t out_xyz[12]) {
    MsmPlan P; P.c = c; P.nwin = nwin;
    msm_combine_host<C>((const XYZZ<C>*)win, P, out_xyz);
}
template void msm_horner_host<0>(const void*, int, int, uint64_t*);
template void msm_horner_host<1>(const void*, int, int, uint64_t*);

template int msm_device<0>(const void*, const void*, size_t, size_t, uint64_t*, ptrdiff_t);
template int msm_device<1>(const void*, const void*, size_t, size_t, uint64_t*, ptrdiff_t);

}  // namespace halo
