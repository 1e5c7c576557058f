// C ABI of libhalo_b200 (include/halo_b200.h): runtime, resident public parameters, MSM entry points,
// arithmetic micro-benchmarks.  Each entry point cites the reference interface it stands in for in the header.
#include "runtime.cuh"
#include <cstring>
#include <algorithm>
#include <thread>
#include <new>
#include <stdexcept>
#if defined(__x86_64__)
#include <emmintrin.h>
#endif

namespace halo {

static Runtime& main_rt() { static Runtime r; return r; }
static thread_local Runtime* t_lane = nullptr;
static Runtime* g_lane = nullptr;
Runtime& rt() { return t_lane ? *t_lane : main_rt(); }
LaneScope::LaneScope(Runtime* r) { t_lane = r; }
LaneScope::~LaneScope() { t_lane = nullptr; }

Runtime* lane_get() {
    Runtime& M = main_rt();
    if (!M.inited) return nullptr;
    if (!g_lane) {
        Runtime* L = new (std::nothrow) Runtime();
        if (!L) return nullptr;
        L->is_lane = true;
        L->device = M.device;
        L->sm_count = M.sm_count;
        bool ok = cudaStreamCreateWithFlags(&L->own_stream, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaStreamCreateWithFlags(&L->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
                  cudaEventCreateWithFlags(&L->copy_done, cudaEventDisableTiming) == cudaSuccess &&
                  cudaEventCreateWithFlags(&L->copy_go, cudaEventDisableTiming) == cudaSuccess;
        if (!ok) { cudaGetLastError(); delete L; return nullptr; }
        L->stream = L->own_stream;
        L->arena.resize(WS_COUNT);
        L->inited = true;
        g_lane = L;
    }
    Runtime* L = g_lane;
    for (int c = 0; c < 2; ++c) L->pp[c] = M.pp[c];
    for (int f = 0; f < 2; ++f) { L->ntt_tw[f] = M.ntt_tw[f]; L->ntt_logmax[f] = M.ntt_logmax[f]; }
    for (int i = 0; i < 16; ++i) L->tune[i] = M.tune[i];
    L->msm_window_override = M.msm_window_override;
    return L;
}
void lane_finish(Runtime* lane) {
    Runtime& M = main_rt();
    M.launches += lane->launches;
    lane->launches = 0;
}
// releases the lane's own resources (the resident tables it points to belong to the main runtime)
static void lane_release() {
    Runtime* L = g_lane;
    if (!L) return;
    cudaStreamSynchronize(L->stream);
    for (auto& b : L->arena) if (b.p) cudaFree(b.p);
    if (L->pinned) cudaFreeHost(L->pinned);
    for (int t = 0; t < Runtime::UP_THREADS; ++t) {
        for (int k = 0; k < Runtime::UP_SLOTS; ++k) {
            if (L->up_buf[t][k]) cudaFreeHost(L->up_buf[t][k]);
            if (L->up_ev[t][k]) cudaEventDestroy(L->up_ev[t][k]);
        }
        if (L->up_stream[t]) cudaStreamDestroy(L->up_stream[t]);
    }
    cudaStreamDestroy(L->own_stream);
    cudaStreamDestroy(L->copy_stream);
    cudaEventDestroy(L->copy_done);
    cudaEventDestroy(L->copy_go);
    for (auto& ev : L->chunk_ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : L->prof_ev) if (ev) cudaEventDestroy(ev);
    delete L;
    g_lane = nullptr;
}

int ws_get(int slot, size_t bytes, void** out) {
    Runtime& R = rt();
    if ((int)R.arena.size() < WS_COUNT) R.arena.resize(WS_COUNT);
    DeviceBuf& b = R.arena[slot];
    if (b.cap < bytes) {
        if (b.p) { HALO_CUDA(cudaStreamSynchronize(R.stream)); HALO_CUDA(cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
        size_t cap = bytes + bytes / 8 + 256;
        HALO_CUDA(cudaMalloc(&b.p, cap));
        b.cap = cap;
    }
    *out = b.p;
    return HALO_OK;
}
int pinned_get(size_t bytes, void** out) {
    Runtime& R = rt();
    if (R.pinned_cap < bytes) {
        if (R.pinned) { HALO_CUDA(cudaStreamSynchronize(R.stream)); HALO_CUDA(cudaFreeHost(R.pinned)); R.pinned = nullptr; R.pinned_cap = 0; }
        size_t cap = bytes < 4096 ? 4096 : bytes;
        HALO_CUDA(cudaMallocHost(&R.pinned, cap));
        R.pinned_cap = cap;
    }
    *out = R.pinned;
    return HALO_OK;
}

bool host_ptr_is_pageable(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

bool dist_on();
int dist_world();
// caller's pageable memory -> pinned chunk.  The chunk is read next by the DMA engine, not by this core: non-temporal
// stores skip the read-for-ownership of the destination lines (a third of the memory traffic of a plain memcpy at these
// sizes, below glibc's own non-temporal threshold) and keep the staging threads from evicting each other's source data.
static inline void stage_copy(void* dst, const void* src, size_t len) {
#if defined(__x86_64__) && defined(__SSE2__)
    if (((uintptr_t)dst & 15) == 0 && len >= 256) {
        const char* s = (const char*)src;
        char* d = (char*)dst;
        size_t i = 0;
        for (; i + 64 <= len; i += 64) {
            const __m128i a = _mm_loadu_si128((const __m128i*)(s + i)), b = _mm_loadu_si128((const __m128i*)(s + i + 16));
            const __m128i c = _mm_loadu_si128((const __m128i*)(s + i + 32)), e = _mm_loadu_si128((const __m128i*)(s + i + 48));
            _mm_stream_si128((__m128i*)(d + i), a); _mm_stream_si128((__m128i*)(d + i + 16), b);
            _mm_stream_si128((__m128i*)(d + i + 32), c); _mm_stream_si128((__m128i*)(d + i + 48), e);
        }
        _mm_sfence();
        if (i < len) memcpy(d + i, s + i, len - i);
        return;
    }
#endif
    memcpy(dst, src, len);
}
int upload_staged(void* d_dst, const void* h_src, size_t bytes, cudaStream_t order_stream) {
    Runtime& R = rt();
    if (!order_stream) order_stream = R.stream;
    if (bytes < 4 * Runtime::UP_CHUNK) {
        HALO_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, order_stream));
        return HALO_OK;
    }
    constexpr int TMAX = Runtime::UP_THREADS, SL = Runtime::UP_SLOTS;
    constexpr size_t CH = Runtime::UP_CHUNK;
    // up to 16 staging threads (2^20-gate columns: 36 -> 30 ms with 16 instead of 8); the ranks of one box share the
    // host's cores, so a process that has joined a communicator takes its share of them
    const int hw = (int)std::max(2u, std::thread::hardware_concurrency());
    const size_t nchunks_all = (bytes + CH - 1) / CH;
    // ... and no more threads than half the chunks: an 8 MB slice of a pipelined MSM upload is four chunks, and starting
    // sixteen threads for it costs more than the copy
    const int T = (int)std::max<size_t>(1, std::min<size_t>((size_t)std::max(2, std::min(TMAX, dist_on() ? hw / std::max(1, dist_world()) : hw)),
                                                            nchunks_all / 2));
    if (!R.up_stream[0]) {
        for (int t = 0; t < TMAX; ++t) {
            HALO_CUDA(cudaStreamCreateWithFlags(&R.up_stream[t], cudaStreamNonBlocking));
            for (int k = 0; k < SL; ++k) {
                HALO_CUDA(cudaMallocHost(&R.up_buf[t][k], CH));
                HALO_CUDA(cudaEventCreateWithFlags(&R.up_ev[t][k], cudaEventDisableTiming));
            }
        }
    }
    // the destination may still be read by queued work: start after the library stream's current tail
    cudaEvent_t go;
    HALO_CUDA(cudaEventCreateWithFlags(&go, cudaEventDisableTiming));
    HALO_CUDA(cudaEventRecord(go, order_stream));          // (the library stream unless the caller orders the copy elsewhere)
    int dev = 0;
    HALO_CUDA(cudaGetDevice(&dev));
    const size_t nchunks = (bytes + CH - 1) / CH;
    int status[TMAX];
    for (int t = 0; t < TMAX; ++t) status[t] = 0;
    Workers wk;
    wk.spawn_indexed(0, (unsigned)T, [&](unsigned t) {
        {
            if (cudaSetDevice(dev) != cudaSuccess || cudaStreamWaitEvent(R.up_stream[t], go, 0) != cudaSuccess) { status[t] = 1; return; }
            int slot = 0;
            for (size_t c = t; c < nchunks; c += T, slot ^= 1) {
                const size_t off = c * CH, len = std::min(CH, bytes - off);
                if (cudaEventSynchronize(R.up_ev[t][slot]) != cudaSuccess) { status[t] = 1; return; }   // slot free again
                // an all-zero chunk (unused columns of a circuit) is set on the device instead of crossing the link;
                // the scan stops at the first nonzero word, so dense data pays nothing for it
                const uint64_t* w = (const uint64_t*)((const char*)h_src + off);
                const size_t words = ((uintptr_t)w % 8 == 0) ? len / 8 : 0;
                size_t nz = 0;
                for (; nz + 64 <= words; nz += 64) {        // 512-byte blocks: vectorisable OR, one branch per block
                    uint64_t any = 0;
                    for (int q = 0; q < 64; ++q) any |= w[nz + q];
                    if (any) break;
                }
                if (nz + 64 > words) while (nz < words && w[nz] == 0) ++nz;
                if (words && nz == words && len % 8 == 0) {
                    if (cudaMemsetAsync((char*)d_dst + off, 0, len, R.up_stream[t]) != cudaSuccess ||
                        cudaEventRecord(R.up_ev[t][slot], R.up_stream[t]) != cudaSuccess) { status[t] = 1; return; }
                    continue;
                }
                stage_copy(R.up_buf[t][slot], (const char*)h_src + off, len);
                if (cudaMemcpyAsync((char*)d_dst + off, R.up_buf[t][slot], len, cudaMemcpyHostToDevice, R.up_stream[t]) != cudaSuccess ||
                    cudaEventRecord(R.up_ev[t][slot], R.up_stream[t]) != cudaSuccess) { status[t] = 1; return; }
            }
        }
    });
    wk.join();
    cudaEventDestroy(go);
    for (int t = 0; t < T; ++t) {
        HALO_CHECK(status[t] == 0, HALO_ECUDA, "staged upload failed");
        for (int k = 0; k < SL; ++k) HALO_CUDA(cudaStreamWaitEvent(order_stream, R.up_ev[t][k], 0));
    }
    return HALO_OK;
}

int msm_wait_bases(cudaStream_t st) {
    Runtime& R = rt();
    if (R.upload_thread.joinable()) {
        R.upload_thread.join();
        HALO_CHECK(R.upload_rc == HALO_OK, HALO_ECUDA, "msm: staged upload of the base points failed");
    }
    if (R.wait_copy) { HALO_CUDA(cudaStreamWaitEvent(st, R.copy_done, 0)); R.wait_copy = false; }
    if (R.bases_hook) {
        auto h = R.bases_hook;
        if (int e = h(st)) return e;
    }
    return HALO_OK;
}
void msm_upload_finish() {
    Runtime& R = rt();
    if (R.upload_thread.joinable()) R.upload_thread.join();
    if (R.wait_copy) { cudaStreamSynchronize(R.copy_stream); R.wait_copy = false; }   // error paths: never leave a copy in flight
}
int msm_upload_inputs(const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k, void** d_s, void** d_b) {
    Runtime& R = rt();
    *d_s = nullptr; *d_b = nullptr;
    if (!k) return HALO_OK;
    if (int e = ws_get(WS_SCALARS, k * 32, d_s)) return e;
    const bool staged = k * 32 >= 4 * Runtime::UP_CHUNK && host_ptr_is_pageable(scalars);
    if (staged) { if (int e = upload_staged(*d_s, scalars, k * 32)) return e; }
    else HALO_CUDA(cudaMemcpyAsync(*d_s, scalars, k * 32, cudaMemcpyHostToDevice, R.stream));
    if (!bases_xy) return HALO_OK;
    // the bases go up on a second stream while the scalar half of the pipeline (digits, scan, scatter, task list)
    // runs; only the accumulation waits for them.  The previous call has finished (MSM calls are synchronous), so
    // the workspace is free to overwrite.
    if (int e = ws_get(WS_BASES, k * 64, d_b)) return e;
    HALO_CUDA(cudaEventRecord(R.copy_go, R.stream));                 // scalars first: they gate the pipeline
    HALO_CUDA(cudaStreamWaitEvent(R.copy_stream, R.copy_go, 0));
    const uint64_t* src = bases_xy + 8 * base_off;
    if (k * 64 >= 4 * Runtime::UP_CHUNK && host_ptr_is_pageable(src)) {
        void* dst = *d_b;
        R.upload_rc = HALO_OK;
        R.wait_copy = true;
        try {
            R.upload_thread = std::thread([dst, src, k]() {
                Runtime& Q = rt();
                int rc = cudaSetDevice(Q.device) == cudaSuccess ? HALO_OK : HALO_ECUDA;
                if (!rc) rc = upload_staged(dst, src, k * 64, Q.copy_stream);
                if (!rc && cudaEventRecord(Q.copy_done, Q.copy_stream) != cudaSuccess) rc = HALO_ECUDA;
                Q.upload_rc = rc;
            });
        } catch (...) {                                              // no thread to be had: do it inline
            if (int e = upload_staged(dst, src, k * 64, R.copy_stream)) return e;
            HALO_CUDA(cudaEventRecord(R.copy_done, R.copy_stream));
        }
    } else {
        HALO_CUDA(cudaMemcpyAsync(*d_b, src, k * 64, cudaMemcpyHostToDevice, R.copy_stream));
        HALO_CUDA(cudaEventRecord(R.copy_done, R.copy_stream));
        R.wait_copy = true;
    }
    return HALO_OK;
}

int prof_begin() {
    Runtime& R = rt();
    R.prof_n = 0;
    if (!R.prof_on) return HALO_OK;
    if (!R.prof_ev[0]) for (int i = 0; i <= Runtime::PROF_MAX; ++i) HALO_CUDA(cudaEventCreate(&R.prof_ev[i]));
    HALO_CUDA(cudaEventRecord(R.prof_ev[0], R.stream));
    return HALO_OK;
}
int prof_mark(const char* name) {
    Runtime& R = rt();
    if (!R.prof_on || R.prof_n >= Runtime::PROF_MAX) return HALO_OK;
    R.prof_name[R.prof_n] = name;
    HALO_CUDA(cudaEventRecord(R.prof_ev[++R.prof_n], R.stream));
    return HALO_OK;
}

template <int C> int msm_device(const void* d_bases, const void* d_scalars, size_t k, size_t nbatch, uint64_t* out_xyz,
                                ptrdiff_t base_stride);
template <int C> int msm_host_chunked(const uint64_t* h_bases, const uint64_t* h_scalars, size_t k, uint64_t* out_xyz);
template <int F> int ntt_device(Fe<F>* data, Fe<F>* out, int log_n, size_t batch, bool inverse, const Fe<F>* coset);
template <int F> int ntt_fold_device(const Fe<F>* coeffs, size_t len, int log_n, Fe<F>* out);
template <int C> int pcdl_commit(const uint64_t*, size_t, size_t, const uint64_t*, uint64_t*);
template <int C> int pcdl_chunked_commit(const uint64_t*, size_t, size_t, const uint64_t*, size_t, uint64_t*, size_t*);
template <int C> int fold_table_build(size_t);
template <int C> int pcdl_open(const uint64_t*, size_t, const uint64_t*, size_t, const uint64_t*, const uint64_t*, const uint64_t*,
                               const uint64_t*, const uint64_t*, uint64_t*, uint64_t*, bool, bool);
template <int C> int pcdl_succinct_check(const uint64_t*, uint64_t*, uint64_t*);
template <int C> int pcdl_check(const uint64_t*);
template <int C> int acc_prover(const uint64_t*, size_t, uint64_t*, bool);
size_t plonk_proof_words(unsigned lg_n);
template <int C> int plonk_prove(unsigned, const uint64_t*, const uint64_t*, const uint64_t*, size_t, const uint64_t*, uint64_t*, const void*, bool, uint64_t*);
int plonk_circuit_shape(int handle, int* curve, unsigned* lg_n);
static inline size_t instance_words(unsigned lg_n) { return 24 * (size_t)lg_n + 54; }
int plonk_circuit_load(int curve, unsigned lg_n, const uint64_t* polys, int* handle);
int plonk_circuit_free(int handle);
void plonk_circuits_release();
int plonk_prove_resident(int handle, const uint64_t*, const uint64_t*, const uint64_t*, size_t, const uint64_t*, uint64_t*, bool);
template <int C> int plonk_verify(unsigned, const uint64_t*, const uint64_t*, size_t, size_t, const uint64_t*, const uint64_t*, int);
template <int C> int acc_verifier(const uint64_t*, size_t, const uint64_t*, bool sharded = false);
template <int C> int acc_prover_dist(const uint64_t*, size_t, uint64_t*);
template <int C> int pcdl_check_dist(const uint64_t*);
bool dist_on();
int dist_rank();
int dist_world();
int dist_allgather_dev(const void* d_send, void* d_recv, size_t bytes);
int dist_bcast_host(void* h_buf, size_t bytes, int root);
template <int S> int hpoly_coeffs(const uint64_t*, const uint64_t*, size_t, unsigned, uint64_t*);
template <int C> int hpoly_commit_words(const uint64_t*, unsigned, uint64_t*);
template <int S> void hpoly_eval_words(const uint64_t*, unsigned, const uint64_t*, uint64_t*);
template <int S> int scalar_dot_host(const uint64_t*, const uint64_t*, size_t, uint64_t*);
template <int S> int construct_powers_host(const uint64_t*, size_t, uint64_t*);
template <int F> int poly_mul_host(const uint64_t*, size_t, const uint64_t*, size_t, uint64_t*);

// ---- synthetic inputs: P_i = k_i * G with k_i = splitmix64(seed, i) | 1  (distinct with overwhelming probability) ----
__host__ __device__ inline u64 splitmix64(u64 x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
template <int C>
__global__ void __launch_bounds__(128) gen_points_kernel(u64 seed, size_t first, size_t n, Affine<C>* out) {
    constexpr int B = CurveTraits<C>::BASE;
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 k = splitmix64(seed + (first + i) * 0x2545F4914F6CDD1DULL) | 1ULL;
    // both Pasta curves use the generator (-1, 2)
    Affine<C> G;
    G.x = fe_neg<B>(fe_one<B>());
    G.y = fe_dbl<B>(fe_one<B>());
    XYZZ<C> acc = xyzz_identity<C>();
    for (int b = 63; b >= 0; --b) {
        acc = xyzz_dbl<C>(acc);
        if ((k >> b) & 1) xyzz_madd<C>(acc, G, false);
    }
    out[i] = xyzz_to_affine<C>(acc);
}

// ---- on-curve check for uploaded parameters (wrappers.rs:606 asserts per point) ----
template <int C>
__global__ void check_on_curve(const Affine<C>* pts, size_t n, int* bad) {
    constexpr int B = CurveTraits<C>::BASE;
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    Affine<C> p = pts[i];
    if (affine_is_identity<C>(p)) return;
    Fe<B> lhs = fe_sqr<B>(p.y);
    Fe<B> rhs = fe_add<B>(fe_mul<B>(fe_sqr<B>(p.x), p.x), fe_from_u64<B>(5));
    if (!fe_eq<B>(lhs, rhs)) atomicAdd(bad, 1);
}

template <int F>
__global__ void field_op_kernel(int op, const Fe<F>* a, const Fe<F>* b, Fe<F>* o, size_t n) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fe<F> x = a[i], y = b ? b[i] : fe_zero<F>(), r;
    switch (op) {
        case 0: r = fe_mul<F>(x, y); break;
        case 1: r = fe_add<F>(x, y); break;
        case 2: r = fe_sub<F>(x, y); break;
        case 3: r = fe_inv<F>(x); break;
        case 4: r = fe_to_canonical<F>(x); break;
        default: r = fe_from_canonical<F>(x); break;
    }
    o[i] = r;
}

// ---- SASS probes: one multiplication / one squaring per thread and nothing else, so that `cuobjdump -sass` shows the
// instruction mix of fe_mul / fe_sqr in isolation (tools/sass_histogram.py -> profiles/).  Never launched.
template <int F> __global__ void sass_probe_mul(const Fe<F>* a, const Fe<F>* b, Fe<F>* o) {
    o[threadIdx.x] = fe_mul<F>(a[threadIdx.x], b[threadIdx.x]);
}
template <int F> __global__ void sass_probe_sqr(const Fe<F>* a, Fe<F>* o) {
    o[threadIdx.x] = fe_sqr<F>(a[threadIdx.x]);
}
template __global__ void sass_probe_mul<1>(const Fe<1>*, const Fe<1>*, Fe<1>*);
template __global__ void sass_probe_sqr<1>(const Fe<1>*, Fe<1>*);

// ---- arithmetic micro-benchmarks ----
template <int F>
__global__ void bench_modmul_kernel(int iters, Fe<F>* out) {
    u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    Fe<F> a[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[j].l[i] = tid * 2654435761u + 40503u * (i + 8 * j) + 17u;
        a[j].l[7] &= 0x3fffffffu;
    }
    Fe<F> m = a[3];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 4; ++j) a[j] = fe_mul<F>(a[j], m);
    }
    Fe<F> s = fe_add<F>(fe_add<F>(a[0], a[1]), fe_add<F>(a[2], a[3]));
    if (s.l[0] == 0x12345678u && s.l[5] == 77u) out[tid & 1023] = s;   // practically never: keeps the work alive
    if (tid == 0) out[0] = s;
}
__global__ void bench_imad_kernel(int iters, unsigned long long* out) {
    u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long acc[8];
    u32 x = tid * 2654435761u + 12345u, y = tid ^ 0x9e3779b9u;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = tid + j;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = (unsigned long long)x * (u32)(y + j) + acc[j];
        x += (u32)acc[0];
    }
    unsigned long long s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s ^= acc[j];
    if (s == 0x123456789abcdefULL) out[tid & 1023] = s;
    if (tid == 0) out[0] = s;
}

// issue-rate probes for the pipes a field multiplication could use: mode 1 = DFMA, 2 = DFMA and IMAD.WIDE interleaved
// 1:1, 3 = 32-bit IMAD, 4 = IADD3 with carry chains
__global__ void bench_pipe_kernel(int mode, int iters, unsigned long long* out) {
    u32 tid = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long acc[8];
    double f[8];
    u32 x = tid * 2654435761u + 12345u, y = tid ^ 0x9e3779b9u;
    double fx = 1.0 + tid * 1e-9, fy = 0.999999 + tid * 1e-12;
#pragma unroll
    for (int j = 0; j < 8; ++j) { acc[j] = tid + j; f[j] = 1.0 + j; }
    if (mode == 1) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = __fma_rz(fx, f[j], fy);
        }
    } else if (mode == 2) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                f[j] = __fma_rz(fx, f[j], fy);
                acc[j] = (unsigned long long)x * (u32)(y + j) + acc[j];
            }
            x += (u32)acc[0];
        }
    } else if (mode == 3) {
        u32 a[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = tid + j;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = a[j] * x + y;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] ^= a[j];
    } else {
        u32 a[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = tid + j;
        for (int it = 0; it < iters; ++it) {
            asm volatile("add.cc.u32 %0, %0, %8; addc.cc.u32 %1, %1, %9; addc.cc.u32 %2, %2, %8; addc.cc.u32 %3, %3, %9;"
                         "addc.cc.u32 %4, %4, %8; addc.cc.u32 %5, %5, %9; addc.cc.u32 %6, %6, %8; addc.u32 %7, %7, %9;"
                         : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7])
                         : "r"(x), "r"(y));
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] ^= a[j];
    }
    unsigned long long s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s ^= acc[j] ^ (unsigned long long)__double_as_longlong(f[j]);
    if (s == 0x123456789abcdefULL) out[tid & 1023] = s;
    if (tid == 0) out[0] = s;
}

}  // namespace halo

using namespace halo;

#define LOCK() std::lock_guard<std::recursive_mutex> lock__(rt().mu)
// Nothing may unwind through the C ABI: std::bad_alloc from a host vector or std::system_error from a thread or mutex
// becomes an error code (the header promises that no entry point aborts the host process).
#define HALO_API_BEGIN try {
#define HALO_API_END } catch (const std::bad_alloc&) { halo::last_err() = "out of host memory"; return HALO_EINVAL; } \
    catch (const std::exception& e__) { halo::last_err() = std::string("exception: ") + e__.what(); return HALO_ECUDA; } \
    catch (...) { halo::last_err() = "unknown exception"; return HALO_ECUDA; }
#define NEED_INIT() HALO_CHECK(rt().inited, HALO_ENOINIT, "halo_init has not been called")

extern "C" int halo_msm(int, const uint64_t*, size_t, const uint64_t*, size_t, uint64_t*);
extern "C" int halo_msm_batch_dev(int, const void*, size_t, const void*, size_t, size_t, uint64_t*);

// point_dot (group.rs:53-56): CurveGroup::normalize_batch (one inversion, Montgomery's trick) on the host — the only
// caller passes m <= 1024 accumulated U's (acc.rs:166) — then the affine MSM on the device.
template <int C>
static int msm_proj_impl(const uint64_t* bases_xyz, const uint64_t* scalars, size_t k, uint64_t out_xyz[12]) {
    constexpr int B = CurveTraits<C>::BASE;
    std::vector<Affine<C>> aff(k);
    std::vector<Fe<B>> pre(k);
    Fe<B> acc = fe_one<B>();
    for (size_t i = 0; i < k; ++i) {
        Jac<C> p; memcpy(&p, bases_xyz + 12 * i, 96);
        pre[i] = acc;
        if (!fe_is_zero<B>(p.Z)) acc = fe_mul<B>(acc, p.Z);
    }
    Fe<B> inv = fe_inv<B>(acc);
    for (size_t i = k; i-- > 0;) {
        Jac<C> p; memcpy(&p, bases_xyz + 12 * i, 96);
        if (fe_is_zero<B>(p.Z)) { aff[i].x = fe_zero<B>(); aff[i].y = fe_zero<B>(); continue; }
        Fe<B> zi = fe_mul<B>(inv, pre[i]);
        inv = fe_mul<B>(inv, p.Z);
        Fe<B> zi2 = fe_sqr<B>(zi);
        aff[i].x = fe_mul<B>(p.X, zi2);
        aff[i].y = fe_mul<B>(p.Y, fe_mul<B>(zi2, zi));
    }
    return halo_msm(C, (const uint64_t*)aff.data(), 0, scalars, k, out_xyz);
}

// sum of n Jacobian points on the host: the combine step of a point-range-sharded MSM (n = number of GPUs)
template <int C> static void point_sum_impl(const uint64_t* pts, size_t n, uint64_t* out) {
    XYZZ<C> acc = xyzz_identity<C>();
    for (size_t i = 0; i < n; ++i) { Jac<C> j; memcpy(&j, pts + 12 * i, 96); xyzz_add<C>(acc, jac_to_xyzz<C>(j)); }
    Jac<C> r = xyzz_to_jac<C>(acc);
    memcpy(out, &r, 96);
}

extern "C" {

const char* halo_last_error(void) { return last_err().c_str(); }

int halo_init(int device) { HALO_API_BEGIN
    LOCK();
    Runtime& R = rt();
    if (R.inited) return HALO_OK;
    if (device >= 0) HALO_CUDA(cudaSetDevice(device));
    HALO_CUDA(cudaGetDevice(&R.device));
    cudaDeviceProp prop;
    HALO_CUDA(cudaGetDeviceProperties(&prop, R.device));
    R.sm_count = prop.multiProcessorCount;
    HALO_CUDA(cudaStreamCreateWithFlags(&R.own_stream, cudaStreamNonBlocking));
    HALO_CUDA(cudaStreamCreateWithFlags(&R.copy_stream, cudaStreamNonBlocking));
    HALO_CUDA(cudaEventCreateWithFlags(&R.copy_done, cudaEventDisableTiming));
    HALO_CUDA(cudaEventCreateWithFlags(&R.copy_go, cudaEventDisableTiming));
    R.stream = R.own_stream;
    R.arena.resize(WS_COUNT);
    R.inited = true;
    return HALO_OK;
HALO_API_END
}

int halo_shutdown(void) { HALO_API_BEGIN
    LOCK();
    Runtime& R = rt();
    if (!R.inited) return HALO_OK;
    cudaStreamSynchronize(R.stream);
    lane_release();
    for (auto& b : R.arena) if (b.p) cudaFree(b.p);
    R.arena.clear();
    for (int c = 0; c < 2; ++c) {
        if (R.pp[c].d_gs) cudaFree(R.pp[c].d_gs);
        if (R.pp[c].d_fold_tab) cudaFree(R.pp[c].d_fold_tab);
        for (void* t : R.pp[c].d_fold_tab_loc) if (t) cudaFree(t);
        if (R.pp[c].d_fixed_tab) cudaFree(R.pp[c].d_fixed_tab);
        R.pp[c] = ResidentParams();
    }
    if (R.pinned) cudaFreeHost(R.pinned);
    R.pinned = nullptr; R.pinned_cap = 0;
    for (int t = 0; t < Runtime::UP_THREADS; ++t) {
        for (int k = 0; k < Runtime::UP_SLOTS; ++k) {
            if (R.up_buf[t][k]) cudaFreeHost(R.up_buf[t][k]);
            if (R.up_ev[t][k]) cudaEventDestroy(R.up_ev[t][k]);
            R.up_buf[t][k] = nullptr; R.up_ev[t][k] = nullptr;
        }
        if (R.up_stream[t]) cudaStreamDestroy(R.up_stream[t]);
        R.up_stream[t] = nullptr;
    }
    for (int f = 0; f < 2; ++f) { if (R.ntt_tw[f]) cudaFree(R.ntt_tw[f]); R.ntt_tw[f] = nullptr; R.ntt_logmax[f] = 0; }
    plonk_circuits_release();
    cudaStreamDestroy(R.own_stream);
    if (R.copy_stream) cudaStreamDestroy(R.copy_stream);
    if (R.copy_done) cudaEventDestroy(R.copy_done);
    if (R.copy_go) cudaEventDestroy(R.copy_go);
    for (auto& ev : R.chunk_ev) { if (ev) cudaEventDestroy(ev); ev = nullptr; }
    R.copy_stream = nullptr; R.copy_done = nullptr;
    R.own_stream = R.stream = nullptr;
    R.inited = false;
    return HALO_OK;
HALO_API_END
}

int halo_set_stream(void* s) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    Runtime& R = rt();
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    R.stream = s ? (cudaStream_t)s : R.own_stream;
    return HALO_OK;
HALO_API_END
}
int halo_device_synchronize(void) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CUDA(cudaStreamSynchronize(rt().stream));
    return HALO_OK;
HALO_API_END
}

int halo_pp_load(int curve, const uint64_t* gs_xy, size_t n, const uint64_t s_xy[8], const uint64_t h_xy[8]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)");
    HALO_CHECK(gs_xy && s_xy && h_xy && n > 0, HALO_EINVAL, "pp_load: null or empty parameters");
    Runtime& R = rt();
    ResidentParams& P = R.pp[curve];
    if (P.n) {
        HALO_CHECK(n <= P.n, HALO_EPARAMS, "public parameters already resident with a smaller length (set_pp is monotone)");
        return HALO_OK;
    }
    void* d = nullptr;
    HALO_CUDA(cudaMalloc(&d, n * 64));
    cudaError_t e = cudaMemcpyAsync(d, gs_xy, n * 64, cudaMemcpyHostToDevice, R.stream);
    int* d_bad = nullptr;
    if (e == cudaSuccess) e = cudaMalloc(&d_bad, 4);
    if (e == cudaSuccess) e = cudaMemsetAsync(d_bad, 0, 4, R.stream);
    if (e == cudaSuccess) {
        unsigned grid = (unsigned)((n + 255) / 256);
        if (curve == 0) check_on_curve<0><<<grid, 256, 0, R.stream>>>((const Affine<0>*)d, n, d_bad);
        else check_on_curve<1><<<grid, 256, 0, R.stream>>>((const Affine<1>*)d, n, d_bad);
        e = cudaGetLastError();
    }
    int bad = 0;
    if (e == cudaSuccess) e = cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, R.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(R.stream);
    if (d_bad) cudaFree(d_bad);
    if (e != cudaSuccess) { cudaFree(d); last_err() = std::string("pp_load: ") + cudaGetErrorString(e); return HALO_ECUDA; }
    if (bad) { cudaFree(d); last_err() = "pp_load: generator not on the curve"; return HALO_EINVAL; }
    P.d_gs = d; P.n = n;
    memcpy(P.s_xy, s_xy, 64); memcpy(P.h_xy, h_xy, 64);
    return HALO_OK;
HALO_API_END
}
// ---- the reference's parameter blobs (crates/group/.precompute/<curve>/gs-NN.bin, sh.bin):
// bincode 2 `standard()` = varint integers: a byte < 251 is the value, 251 / 252 / 253 introduce a u16 / u32 / u64 LE.
// gs-NN.bin is a Vec<WrappedPoint>: a length, then per point the eight u64 limbs x[0..4], y[0..4] (Montgomery);
// sh.bin is the tuple (S, H) without a length.
static bool read_varint(const uint8_t* b, size_t len, size_t& pos, uint64_t& v) {
    if (pos >= len) return false;
    const uint8_t t = b[pos++];
    if (t < 251) { v = t; return true; }
    const size_t w = t == 251 ? 2 : t == 252 ? 4 : t == 253 ? 8 : 0;
    if (!w || pos + w > len) return false;
    v = 0;
    for (size_t i = 0; i < w; ++i) v |= (uint64_t)b[pos + i] << (8 * i);
    pos += w;
    return true;
}
int halo_pp_decode_block(const uint8_t* bytes, size_t len, int with_length, size_t max_points, uint64_t* out_xy,
                         size_t* n_points) { HALO_API_BEGIN
    HALO_CHECK(bytes && out_xy && n_points, HALO_EINVAL, "pp_decode_block: null argument");
    size_t pos = 0;
    uint64_t count = max_points;
    if (with_length) HALO_CHECK(read_varint(bytes, len, pos, count), HALO_EINVAL, "pp_decode_block: truncated length");
    const size_t take = (size_t)std::min<uint64_t>(count, max_points);           // .into_iter().take(m), pp.rs:41
    for (size_t i = 0; i < take; ++i)
        for (int l = 0; l < 8; ++l)
            HALO_CHECK(read_varint(bytes, len, pos, out_xy[8 * i + l]), HALO_EINVAL, "pp_decode_block: truncated or malformed point");
    *n_points = take;
    return HALO_OK;
HALO_API_END
}
// PublicParams::new(n) (crates/group/src/pp.rs:26-61): the first n generators of the blocks, S and H from sh.bin,
// then resident exactly as halo_pp_load makes them (on-curve check on the device, pp.rs:27-32 / wrappers.rs:603-616)
int halo_pp_load_blobs(int curve, const uint8_t* const* blocks, const size_t* block_lens, size_t n_blocks,
                       const uint8_t* sh, size_t sh_len, size_t n) { HALO_API_BEGIN
    HALO_CHECK(blocks && block_lens && sh && n_blocks > 0, HALO_EINVAL, "pp_load_blobs: null argument");
    HALO_CHECK(n > 0 && (n & (n - 1)) == 0, HALO_EINVAL, "pp_load_blobs: n must be a power of two (pp.rs:30)");
    // every block's length prefix first, so that the blocks decode in parallel into their final positions
    std::vector<size_t> first(n_blocks + 1, 0), take(n_blocks, 0);
    for (size_t i = 0; i < n_blocks && first[i] < n; ++i) {
        size_t pos = 0; uint64_t cnt = 0;
        HALO_CHECK(blocks[i] && read_varint(blocks[i], block_lens[i], pos, cnt), HALO_EINVAL, "pp_load_blobs: failed to decode a block");
        take[i] = (size_t)std::min<uint64_t>(cnt, n - first[i]);
        first[i + 1] = first[i] + take[i];
        for (size_t j = i + 2; j <= n_blocks; ++j) first[j] = first[i + 1];
    }
    HALO_CHECK(first[n_blocks] == n, HALO_EPARAMS, "pp_load_blobs: the blocks hold fewer than n generators (pp.rs:31)");
    std::vector<uint64_t> gs(n * 8);
    std::vector<int> status(n_blocks, HALO_OK);
    {
        const unsigned nt = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)n_blocks));
        Workers wk;
        wk.spawn_indexed(0, nt, [&](unsigned t) {
            for (size_t i = t; i < n_blocks; i += nt) {
                if (!take[i]) continue;
                size_t got = 0;
                status[i] = halo_pp_decode_block(blocks[i], block_lens[i], 1, take[i], gs.data() + 8 * first[i], &got);
                if (status[i] == HALO_OK && got != take[i]) status[i] = HALO_EINVAL;
            }
        });
        wk.join();
    }
    for (size_t i = 0; i < n_blocks; ++i) HALO_CHECK(status[i] == HALO_OK, HALO_EINVAL, "pp_load_blobs: failed to decode a block");
    uint64_t shw[16]; size_t two = 0;
    if (int e = halo_pp_decode_block(sh, sh_len, 0, 2, shw, &two)) return e;
    return halo_pp_load(curve, gs.data(), n, shw, shw + 8);
HALO_API_END
}
size_t halo_pp_len(int curve) {
    try { LOCK(); return (curve == 0 || curve == 1) ? rt().pp[curve].n : 0; } catch (...) { return 0; }
}
int halo_pp_reset(int curve) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    Runtime& R = rt();
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    if (R.pp[curve].d_gs) cudaFree(R.pp[curve].d_gs);
    if (R.pp[curve].d_fold_tab) cudaFree(R.pp[curve].d_fold_tab);
    for (void* t : R.pp[curve].d_fold_tab_loc) if (t) cudaFree(t);
    if (R.pp[curve].d_fixed_tab) cudaFree(R.pp[curve].d_fixed_tab);
    R.pp[curve] = ResidentParams();
    return HALO_OK;
HALO_API_END
}

int halo_profile_enable(int on) { HALO_API_BEGIN LOCK(); rt().prof_on = on != 0; rt().prof_n = 0; return HALO_OK; HALO_API_END }
int halo_profile_read(float* ms, const char** names, int max_stages) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    Runtime& R = rt();
    int n = R.prof_n < max_stages ? R.prof_n : max_stages;
    if (n > 0) HALO_CUDA(cudaEventSynchronize(R.prof_ev[R.prof_n]));
    for (int i = 0; i < n; ++i) {
        HALO_CUDA(cudaEventElapsedTime(&ms[i], R.prof_ev[i], R.prof_ev[i + 1]));
        if (names) names[i] = R.prof_name[i];
    }
    return n;
HALO_API_END
}
unsigned long long halo_launch_count(void) { return rt().launches; }

uint64_t halo_gen_point_multiplier(uint64_t seed, uint64_t i) { return splitmix64(seed + i * 0x2545F4914F6CDD1DULL) | 1ULL; }
int halo_gen_points_dev(int curve, uint64_t seed, size_t first, size_t n, void* d_out_xy) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((curve == 0 || curve == 1) && (d_out_xy || n == 0), HALO_EINVAL, "gen_points: bad argument");
    if (n == 0) return HALO_OK;
    Runtime& R = rt();
    unsigned grid = (unsigned)((n + 127) / 128);
    if (curve == 0) gen_points_kernel<0><<<grid, 128, 0, R.stream>>>(seed, first, n, (Affine<0>*)d_out_xy);
    else gen_points_kernel<1><<<grid, 128, 0, R.stream>>>(seed, first, n, (Affine<1>*)d_out_xy);
    HALO_CUDA(cudaGetLastError());
    return HALO_OK;
HALO_API_END
}
int halo_gen_points(int curve, uint64_t seed, size_t first, size_t n, uint64_t* out_xy) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(out_xy || n == 0, HALO_EINVAL, "gen_points: null output");
    if (n == 0) return HALO_OK;
    void* d;
    if (int e = ws_get(WS_BASES, n * 64, &d)) return e;
    if (int e = halo_gen_points_dev(curve, seed, first, n, d)) return e;
    HALO_CUDA(cudaMemcpyAsync(out_xy, d, n * 64, cudaMemcpyDeviceToHost, rt().stream));
    HALO_CUDA(cudaStreamSynchronize(rt().stream));
    return HALO_OK;
HALO_API_END
}

int halo_set_tuning(int key, int value) { HALO_API_BEGIN LOCK(); if (key < 0 || key >= 16) return HALO_EINVAL; rt().tune[key] = value; return HALO_OK; HALO_API_END }
int halo_msm_set_window(int c) { HALO_API_BEGIN LOCK(); rt().msm_window_override = c; return HALO_OK; HALO_API_END }

int halo_msm_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (d_scalars || k == 0), HALO_EINVAL, "msm: null pointer");
    Runtime& R = rt();
    const char* bases = (const char*)d_bases_xy;
    if (!bases) {
        HALO_CHECK(R.pp[curve].n >= base_off + k, HALO_EPARAMS, "msm: resident generators missing or too few");
        bases = (const char*)R.pp[curve].d_gs;
    }
    bases += base_off * 64;
    return curve == 0 ? msm_device<0>(bases, d_scalars, k, 1, out_xyz, 0) : msm_device<1>(bases, d_scalars, k, 1, out_xyz, 0);
HALO_API_END
}

int halo_msm_batch_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k, size_t b,
                       uint64_t* out_xyz) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (d_scalars || k == 0 || b == 0), HALO_EINVAL, "msm_batch: null pointer");
    Runtime& R = rt();
    const char* bases = (const char*)d_bases_xy;
    if (!bases) {
        HALO_CHECK(R.pp[curve].n >= base_off + k, HALO_EPARAMS, "msm_batch: resident generators missing or too few");
        bases = (const char*)R.pp[curve].d_gs;
    }
    bases += base_off * 64;
    return curve == 0 ? msm_device<0>(bases, d_scalars, k, b, out_xyz, 0) : msm_device<1>(bases, d_scalars, k, b, out_xyz, 0);
HALO_API_END
}

int halo_msm(int curve, const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (scalars || k == 0), HALO_EINVAL, "msm: null pointer");
    // large MSMs over caller-supplied bases: upload pipelined with the accumulation by point range (tuning key 9 = 1: off)
    if (bases_xy && k >= ((size_t)1 << 19) && k < ((size_t)1 << 27) && rt().tune[9] != 1 && rt().tune[8] == 0)
        return curve == 0 ? msm_host_chunked<0>(bases_xy + 8 * base_off, scalars, k, out_xyz)
                          : msm_host_chunked<1>(bases_xy + 8 * base_off, scalars, k, out_xyz);
    void *d_s = nullptr, *d_b = nullptr;
    if (int e = msm_upload_inputs(bases_xy, base_off, scalars, k, &d_s, &d_b)) { msm_upload_finish(); return e; }
    if (d_b) base_off = 0;
    int rc = halo_msm_dev(curve, d_b, base_off, d_s, k, out_xyz);
    msm_upload_finish();
    return rc;
HALO_API_END
}

int halo_msm_batch(int curve, size_t base_off, const uint64_t* scalars, size_t k, size_t b, uint64_t* out_xyz) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (scalars || k == 0 || b == 0), HALO_EINVAL, "msm_batch: null pointer");
    void* d_s = nullptr;
    if (k && b) {
        if (int e = ws_get(WS_SCALARS, b * k * 32, &d_s)) return e;
        HALO_CUDA(cudaMemcpyAsync(d_s, scalars, b * k * 32, cudaMemcpyHostToDevice, rt().stream));
    }
    return halo_msm_batch_dev(curve, nullptr, base_off, d_s, k, b, out_xyz);
HALO_API_END
}

int halo_msm_proj(int curve, const uint64_t* bases_xyz, const uint64_t* scalars, size_t k, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && ((bases_xyz && scalars) || k == 0), HALO_EINVAL, "msm_proj: null pointer");
    return curve == 0 ? msm_proj_impl<0>(bases_xyz, scalars, k, out_xyz) : msm_proj_impl<1>(bases_xyz, scalars, k, out_xyz);
HALO_API_END
}

int halo_point_sum(int curve, const uint64_t* pts_xyz, size_t n, uint64_t out_xyz[12]) { HALO_API_BEGIN
    HALO_CHECK((curve == 0 || curve == 1) && out_xyz && (pts_xyz || n == 0), HALO_EINVAL, "point_sum: bad argument");
    if (curve == 0) point_sum_impl<0>(pts_xyz, n, out_xyz); else point_sum_impl<1>(pts_xyz, n, out_xyz);
    return HALO_OK;
HALO_API_END
}

int halo_field_op(int field, int op, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && a && out && op >= 0 && op <= 5, HALO_EINVAL, "field_op: bad argument");
    if (n == 0) return HALO_OK;
    Runtime& R = rt();
    void *da, *db = nullptr, *dout;
    if (int e = ws_get(WS_TMP0, n * 32, &da)) return e;
    if (int e = ws_get(WS_TMP2, n * 32, &dout)) return e;
    HALO_CUDA(cudaMemcpyAsync(da, a, n * 32, cudaMemcpyHostToDevice, R.stream));
    if (b) {
        if (int e = ws_get(WS_TMP1, n * 32, &db)) return e;
        HALO_CUDA(cudaMemcpyAsync(db, b, n * 32, cudaMemcpyHostToDevice, R.stream));
    }
    unsigned grid = (unsigned)((n + 127) / 128);
    if (field == 0) field_op_kernel<0><<<grid, 128, 0, R.stream>>>(op, (const Fe<0>*)da, (const Fe<0>*)db, (Fe<0>*)dout, n);
    else field_op_kernel<1><<<grid, 128, 0, R.stream>>>(op, (const Fe<1>*)da, (const Fe<1>*)db, (Fe<1>*)dout, n);
    HALO_CUDA(cudaGetLastError());
    HALO_CUDA(cudaMemcpyAsync(out, dout, n * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
HALO_API_END
}

int halo_bench_modmul(int field, int blocks, int threads, int iters, float* ms, uint64_t out[4]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    Runtime& R = rt();
    void* d;
    if (int e = ws_get(WS_TMP0, 1024 * 32, &d)) return e;
    cudaEvent_t e0, e1;
    HALO_CUDA(cudaEventCreate(&e0)); HALO_CUDA(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; ++rep) {      // first pass warms up
        HALO_CUDA(cudaEventRecord(e0, R.stream));
        if (field == 0) bench_modmul_kernel<0><<<blocks, threads, 0, R.stream>>>(iters, (Fe<0>*)d);
        else bench_modmul_kernel<1><<<blocks, threads, 0, R.stream>>>(iters, (Fe<1>*)d);
        HALO_CUDA(cudaEventRecord(e1, R.stream));
        HALO_CUDA(cudaStreamSynchronize(R.stream));
    }
    HALO_CUDA(cudaGetLastError());
    HALO_CUDA(cudaEventElapsedTime(ms, e0, e1));
    HALO_CUDA(cudaMemcpy(out, d, 32, cudaMemcpyDeviceToHost));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return HALO_OK;
HALO_API_END
}
int halo_bench_imad(int blocks, int threads, int iters, float* ms, uint64_t* out) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    Runtime& R = rt();
    void* d;
    if (int e = ws_get(WS_TMP0, 1024 * 32, &d)) return e;
    cudaEvent_t e0, e1;
    HALO_CUDA(cudaEventCreate(&e0)); HALO_CUDA(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; ++rep) {
        HALO_CUDA(cudaEventRecord(e0, R.stream));
        bench_imad_kernel<<<blocks, threads, 0, R.stream>>>(iters, (unsigned long long*)d);
        HALO_CUDA(cudaEventRecord(e1, R.stream));
        HALO_CUDA(cudaStreamSynchronize(R.stream));
    }
    HALO_CUDA(cudaGetLastError());
    HALO_CUDA(cudaEventElapsedTime(ms, e0, e1));
    HALO_CUDA(cudaMemcpy(out, d, 8, cudaMemcpyDeviceToHost));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return HALO_OK;
HALO_API_END
}

// ---- PLONK prover / verifier (plonk.cu) ----
size_t halo_plonk_proof_words(unsigned lg_n) { return plonk_proof_words(lg_n); }

int halo_plonk_prove(int curve, unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                     size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)");
    HALO_CHECK(polys && w_evals && acc_prev && out_proof && (public_inputs || n_public == 0), HALO_EINVAL, "null argument");
    return curve == 0 ? plonk_prove<0>(lg_n, polys, w_evals, public_inputs, n_public, acc_prev, out_proof, nullptr, false, nullptr)
                      : plonk_prove<1>(lg_n, polys, w_evals, public_inputs, n_public, acc_prev, out_proof, nullptr, false, nullptr);
HALO_API_END
}

// f_GC on H (the gate constraints of protocol.rs:170-241 without the permutation terms): out[j] = f_GC(omega^j), n x 4
// words.  All zero exactly when the witness satisfies every gate row.
int halo_plonk_gate_residual(int curve, unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                             size_t n_public, uint64_t* out_evals) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)");
    HALO_CHECK(polys && w_evals && out_evals && (public_inputs || n_public == 0), HALO_EINVAL, "null argument");
    HALO_CHECK(lg_n >= 1 && lg_n <= 22, HALO_EINVAL, "plonk: rows must be a power of two in [2, 2^22]");
    std::vector<uint64_t> acc(instance_words(lg_n), 0), proof(plonk_proof_words(lg_n), 0);
    acc[12] = ((uint64_t)1 << lg_n) - 1;
    Runtime& R = rt();
    const int t15 = R.tune[15];
    R.tune[15] = 1;                                        // no helper thread: the run stops before round 1's result is used
    const int rc = curve == 0 ? plonk_prove<0>(lg_n, polys, w_evals, public_inputs, n_public, acc.data(), proof.data(), nullptr, false, out_evals)
                              : plonk_prove<1>(lg_n, polys, w_evals, public_inputs, n_public, acc.data(), proof.data(), nullptr, false, out_evals);
    R.tune[15] = t15;
    return rc;
HALO_API_END
}

int halo_plonk_circuit_load(int curve, unsigned lg_n, const uint64_t* polys, int* handle) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)");
    HALO_CHECK(polys && handle, HALO_EINVAL, "null argument");
    return plonk_circuit_load(curve, lg_n, polys, handle);
HALO_API_END
}
int halo_plonk_circuit_free(int handle) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    return plonk_circuit_free(handle);
HALO_API_END
}
int halo_plonk_prove_circuit(int handle, const uint64_t* w_polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                             size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(w_polys && w_evals && acc_prev && out_proof && (public_inputs || n_public == 0), HALO_EINVAL, "null argument");
    return plonk_prove_resident(handle, w_polys, w_evals, public_inputs, n_public, acc_prev, out_proof, false);
HALO_API_END
}

// ONE proof over the GPUs of the communicator (collective).  The small inputs (public inputs, previous accumulator)
// are broadcast from rank 0 here, the columns inside plonk_prove.
static int dist_plonk_small_inputs(unsigned lg_n, const uint64_t* public_inputs, size_t n_public, const uint64_t* acc_prev,
                                   std::vector<uint64_t>& pub, std::vector<uint64_t>& acc) {
    HALO_CHECK(lg_n >= 1 && lg_n <= 22, HALO_EINVAL, "plonk: rows must be a power of two in [2, 2^22]");
    HALO_CHECK(n_public <= ((size_t)1 << lg_n), HALO_EINVAL, "plonk: more public inputs than rows");
    pub.assign(4 * n_public, 0);
    acc.assign(instance_words(lg_n), 0);
    if (dist_rank() == 0) {
        HALO_CHECK(acc_prev && (public_inputs || n_public == 0), HALO_EINVAL, "null argument on rank 0");
        if (n_public) memcpy(pub.data(), public_inputs, pub.size() * 8);
        memcpy(acc.data(), acc_prev, acc.size() * 8);
    }
    if (int e = dist_bcast_host(pub.data(), pub.size() * 8, 0)) return e;
    return dist_bcast_host(acc.data(), acc.size() * 8, 0);
}
int halo_dist_plonk_prove(int curve, unsigned lg_n, const uint64_t* polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                          size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)");
    HALO_CHECK(out_proof && (dist_rank() != 0 || (polys && w_evals)), HALO_EINVAL, "null argument");
    std::vector<uint64_t> pub, acc;
    if (int e = dist_plonk_small_inputs(lg_n, public_inputs, n_public, acc_prev, pub, acc)) return e;
    return curve == 0 ? plonk_prove<0>(lg_n, polys, w_evals, pub.data(), n_public, acc.data(), out_proof, nullptr, true, nullptr)
                      : plonk_prove<1>(lg_n, polys, w_evals, pub.data(), n_public, acc.data(), out_proof, nullptr, true, nullptr);
HALO_API_END
}
int halo_dist_plonk_prove_circuit(int handle, const uint64_t* w_polys, const uint64_t* w_evals, const uint64_t* public_inputs,
                                  size_t n_public, const uint64_t* acc_prev, uint64_t* out_proof) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    HALO_CHECK(out_proof && (dist_rank() != 0 || (w_polys && w_evals)), HALO_EINVAL, "null argument");
    int curve = 0; unsigned lg_n = 0;
    if (int e = plonk_circuit_shape(handle, &curve, &lg_n)) return e;
    std::vector<uint64_t> pub, acc;
    if (int e = dist_plonk_small_inputs(lg_n, public_inputs, n_public, acc_prev, pub, acc)) return e;
    return plonk_prove_resident(handle, w_polys, w_evals, pub.data(), n_public, acc.data(), out_proof, true);
HALO_API_END
}

int halo_plonk_verify(int curve, unsigned lg_n, const uint64_t* circuit_C_qs, const uint64_t* public_inputs, size_t n_public,
                      size_t public_input_count, const uint64_t* acc_prev, const uint64_t* proof, int succinct_only) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)");
    HALO_CHECK(circuit_C_qs && acc_prev && proof && (public_inputs || n_public == 0), HALO_EINVAL, "null argument");
    return curve == 0 ? plonk_verify<0>(lg_n, circuit_C_qs, public_inputs, n_public, public_input_count, acc_prev, proof, succinct_only)
                      : plonk_verify<1>(lg_n, circuit_C_qs, public_inputs, n_public, public_input_count, acc_prev, proof, succinct_only);
HALO_API_END
}

int halo_bench_pipe(int mode, int blocks, int threads, int iters, float* ms, uint64_t* out) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    Runtime& R = rt();
    void* d;
    if (int e = ws_get(WS_TMP0, 1024 * 32, &d)) return e;
    cudaEvent_t e0, e1;
    HALO_CUDA(cudaEventCreate(&e0)); HALO_CUDA(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; ++rep) {
        HALO_CUDA(cudaEventRecord(e0, R.stream));
        bench_pipe_kernel<<<blocks, threads, 0, R.stream>>>(mode, iters, (unsigned long long*)d);
        HALO_CUDA(cudaEventRecord(e1, R.stream));
        HALO_CUDA(cudaStreamSynchronize(R.stream));
    }
    HALO_CUDA(cudaGetLastError());
    HALO_CUDA(cudaEventElapsedTime(ms, e0, e1));
    HALO_CUDA(cudaMemcpy(out, d, 8, cudaMemcpyDeviceToHost));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return HALO_OK;
HALO_API_END
}

}  // extern "C"

// ---- NTT ----
template <int F>
static int ntt_dev_impl(void* d_data, unsigned log_n, size_t b, int inverse, const uint64_t* coset) {
    Runtime& R = rt();
    const size_t bytes = ((size_t)b << log_n) * 32;
    void* scratch;
    if (int e = ws_get(WS_NTT1, bytes, &scratch)) return e;
    Fe<F> g;
    if (coset) memcpy(&g, coset, 32);
    if (int e = ntt_device<F>((Fe<F>*)d_data, (Fe<F>*)scratch, (int)log_n, b, inverse != 0, coset ? &g : nullptr)) return e;
    HALO_CUDA(cudaMemcpyAsync(d_data, scratch, bytes, cudaMemcpyDeviceToDevice, R.stream));
    return HALO_OK;
}
template <int F>
static int ntt_host_impl(uint64_t* data, unsigned log_n, size_t b, int inverse, const uint64_t* coset) {
    Runtime& R = rt();
    const size_t bytes = ((size_t)b << log_n) * 32;
    void *d_in, *d_out;
    if (int e = ws_get(WS_NTT0, bytes, &d_in)) return e;
    if (int e = ws_get(WS_NTT1, bytes, &d_out)) return e;
    HALO_CUDA(cudaMemcpyAsync(d_in, data, bytes, cudaMemcpyHostToDevice, R.stream));
    Fe<F> g;
    if (coset) memcpy(&g, coset, 32);
    if (int e = ntt_device<F>((Fe<F>*)d_in, (Fe<F>*)d_out, (int)log_n, b, inverse != 0, coset ? &g : nullptr)) return e;
    HALO_CUDA(cudaMemcpyAsync(data, d_out, bytes, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
}
template <int F>
static int poly_evaluate_impl(const uint64_t* coeffs, size_t len, unsigned log_n, uint64_t* out_evals) {
    Runtime& R = rt();
    const size_t N = (size_t)1 << log_n;
    void *d_c, *d_in, *d_out;
    if (int e = ws_get(WS_NTT0, N * 32, &d_in)) return e;
    if (int e = ws_get(WS_NTT1, N * 32, &d_out)) return e;
    if (len) {
        if (int e = ws_get(WS_TMP0, len * 32, &d_c)) return e;
        HALO_CUDA(cudaMemcpyAsync(d_c, coeffs, len * 32, cudaMemcpyHostToDevice, R.stream));
        if (int e = ntt_fold_device<F>((const Fe<F>*)d_c, len, (int)log_n, (Fe<F>*)d_in)) return e;
    } else {
        HALO_CUDA(cudaMemsetAsync(d_in, 0, N * 32, R.stream));
    }
    if (int e = ntt_device<F>((Fe<F>*)d_in, (Fe<F>*)d_out, (int)log_n, 1, false, nullptr)) return e;
    HALO_CUDA(cudaMemcpyAsync(out_evals, d_out, N * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
}

// NTT / iNTT batches across the communicator, by polynomial: rank r transforms the polynomials [r*per, (r+1)*per) of the
// batch (per = ceil(b / world)); the blocks are all-gathered between device buffers and every rank returns the whole
// result.  A single transform never leaves its GPU.
template <int F>
static int ntt_dist_impl(uint64_t* data, unsigned log_n, size_t b, int inverse, const uint64_t* coset) {
    Runtime& R = rt();
    const size_t W = (size_t)dist_world(), me = (size_t)dist_rank();
    const size_t per = (b + W - 1) / W, N = (size_t)1 << log_n, blk = per * N * 32;
    const size_t lo = std::min(b, me * per), hi = std::min(b, lo + per), cnt = hi - lo;
    void *d_in, *d_out, *d_all;
    if (int e = ws_get(WS_NTT0, blk, &d_in)) return e;
    if (int e = ws_get(WS_NTT1, blk, &d_out)) return e;
    if (int e = ws_get(WS_PLONK_TMP, blk * W, &d_all)) return e;
    if (cnt) {
        if (int e = upload_staged(d_in, data + lo * N * 4, cnt * N * 32)) return e;
        Fe<F> g;
        if (coset) memcpy(&g, coset, 32);
        if (int e = ntt_device<F>((Fe<F>*)d_in, (Fe<F>*)d_out, (int)log_n, cnt, inverse != 0, coset ? &g : nullptr)) return e;
    }
    if (cnt < per) HALO_CUDA(cudaMemsetAsync((char*)d_out + cnt * N * 32, 0, (per - cnt) * N * 32, R.stream));
    if (int e = dist_allgather_dev(d_out, d_all, blk)) return e;
    HALO_CUDA(cudaMemcpyAsync(data, d_all, b * N * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
}

extern "C" {
int halo_dist_ntt(int field, uint64_t* data, unsigned log_n, size_t b, int inverse, const uint64_t* coset_offset_or_null) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && (data || b == 0) && log_n <= 30, HALO_EINVAL, "ntt: bad argument");
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    if (b == 0) return HALO_OK;
    return field == 0 ? ntt_dist_impl<0>(data, log_n, b, inverse, coset_offset_or_null)
                      : ntt_dist_impl<1>(data, log_n, b, inverse, coset_offset_or_null);
HALO_API_END
}
int halo_ntt_dev(int field, void* d_data, unsigned log_n, size_t b, int inverse, const uint64_t* coset_offset_or_null) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && (d_data || b == 0) && log_n <= 30, HALO_EINVAL, "ntt: bad argument");
    if (b == 0) return HALO_OK;
    return field == 0 ? ntt_dev_impl<0>(d_data, log_n, b, inverse, coset_offset_or_null)
                      : ntt_dev_impl<1>(d_data, log_n, b, inverse, coset_offset_or_null);
HALO_API_END
}
int halo_ntt(int field, uint64_t* data, unsigned log_n, size_t b, int inverse, const uint64_t* coset_offset_or_null) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && (data || b == 0) && log_n <= 30, HALO_EINVAL, "ntt: bad argument");
    if (b == 0) return HALO_OK;
    return field == 0 ? ntt_host_impl<0>(data, log_n, b, inverse, coset_offset_or_null)
                      : ntt_host_impl<1>(data, log_n, b, inverse, coset_offset_or_null);
HALO_API_END
}
int halo_poly_evaluate(int field, const uint64_t* coeffs, size_t len, unsigned log_n, uint64_t* out_evals) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && (coeffs || len == 0) && out_evals && log_n <= 30, HALO_EINVAL, "poly_evaluate: bad argument");
    return field == 0 ? poly_evaluate_impl<0>(coeffs, len, log_n, out_evals) : poly_evaluate_impl<1>(coeffs, len, log_n, out_evals);
HALO_API_END
}
}  // extern "C"

// ---- PC_DL / AS_DL ----
#define NEED_CURVE() HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "curve must be 0 (Pallas) or 1 (Vesta)")
extern "C" {
size_t halo_instance_words(unsigned lg_n) { return 24 * (size_t)lg_n + 54; }
size_t halo_proof_words(unsigned lg_n) { return 24 * (size_t)lg_n + 33; }

int halo_pcdl_commit(int curve, const uint64_t* coeffs, size_t len, size_t d, const uint64_t* w_or_null, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(out_xyz && (coeffs || len == 0), HALO_EINVAL, "pcdl_commit: null pointer");
    return curve == 0 ? pcdl_commit<0>(coeffs, len, d, w_or_null, out_xyz) : pcdl_commit<1>(coeffs, len, d, w_or_null, out_xyz);
HALO_API_END
}
int halo_pcdl_chunked_commit(int curve, const uint64_t* coeffs, size_t len, size_t d, const uint64_t* w_or_null, size_t chunk_size,
                              uint64_t* out_xyz, size_t* out_count) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK((out_xyz || len == 0) && (coeffs || len == 0), HALO_EINVAL, "chunked_commit: null pointer");
    return curve == 0 ? pcdl_chunked_commit<0>(coeffs, len, d, w_or_null, chunk_size, out_xyz, out_count)
                      : pcdl_chunked_commit<1>(coeffs, len, d, w_or_null, chunk_size, out_xyz, out_count);
HALO_API_END
}
int halo_ipa_fold_table_build(int curve, size_t n) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    return curve == 0 ? fold_table_build<0>(n) : fold_table_build<1>(n);
HALO_API_END
}
int halo_pcdl_open(int curve, const uint64_t* coeffs, size_t len, const uint64_t C_xyz[12], size_t d, const uint64_t z[4],
                   const uint64_t* v_or_null, const uint64_t* w_or_null, const uint64_t* hide_q, const uint64_t* hide_wbar,
                   uint64_t* out_proof, uint64_t* out_v_or_null) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(C_xyz && z && out_proof && (coeffs || len == 0), HALO_EINVAL, "pcdl_open: null pointer");
    return curve == 0 ? pcdl_open<0>(coeffs, len, C_xyz, d, z, v_or_null, w_or_null, hide_q, hide_wbar, out_proof, out_v_or_null, false, false)
                      : pcdl_open<1>(coeffs, len, C_xyz, d, z, v_or_null, w_or_null, hide_q, hide_wbar, out_proof, out_v_or_null, false, false);
HALO_API_END
}
// pcdl::open over the GPUs of the communicator (collective, the same inputs on every rank, the same proof on every rank):
// the rounds above the fold-free tail run on interleaved slices of the generators (ipa.cu, pcdl_open with shard = true)
int halo_dist_pcdl_open(int curve, const uint64_t* coeffs, size_t len, const uint64_t C_xyz[12], size_t d, const uint64_t z[4],
                        const uint64_t* v_or_null, const uint64_t* w_or_null, const uint64_t* hide_q, const uint64_t* hide_wbar,
                        uint64_t* out_proof, uint64_t* out_v_or_null) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    HALO_CHECK(C_xyz && z && out_proof && (coeffs || len == 0), HALO_EINVAL, "pcdl_open: null pointer");
    return curve == 0 ? pcdl_open<0>(coeffs, len, C_xyz, d, z, v_or_null, w_or_null, hide_q, hide_wbar, out_proof, out_v_or_null, false, true)
                      : pcdl_open<1>(coeffs, len, C_xyz, d, z, v_or_null, w_or_null, hide_q, hide_wbar, out_proof, out_v_or_null, false, true);
HALO_API_END
}
int halo_pcdl_succinct_check(int curve, const uint64_t* instance, uint64_t* out_xis_or_null, uint64_t* out_U_or_null) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(instance, HALO_EINVAL, "succinct_check: null pointer");
    return curve == 0 ? pcdl_succinct_check<0>(instance, out_xis_or_null, out_U_or_null)
                      : pcdl_succinct_check<1>(instance, out_xis_or_null, out_U_or_null);
HALO_API_END
}
int halo_pcdl_check(int curve, const uint64_t* instance) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(instance, HALO_EINVAL, "check: null pointer");
    return curve == 0 ? pcdl_check<0>(instance) : pcdl_check<1>(instance);
HALO_API_END
}
int halo_acc_prover(int curve, const uint64_t* instances, size_t m, uint64_t* out_acc) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(out_acc, HALO_EINVAL, "acc_prover: null pointer");
    return curve == 0 ? acc_prover<0>(instances, m, out_acc, false) : acc_prover<1>(instances, m, out_acc, false);
HALO_API_END
}
int halo_acc_verifier(int curve, const uint64_t* instances, size_t m, const uint64_t* acc) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(acc, HALO_EINVAL, "acc_verifier: null pointer");
    return curve == 0 ? acc_verifier<0>(instances, m, acc) : acc_verifier<1>(instances, m, acc);
HALO_API_END
}
int halo_acc_decider(int curve, const uint64_t* acc) { HALO_API_BEGIN return halo_pcdl_check(curve, acc); HALO_API_END }
/* ---- the same three across the GPUs of the library's communicator (dist.cu) ---- */
int halo_dist_acc_prover(int curve, const uint64_t* instances, size_t m, uint64_t* out_acc) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(out_acc, HALO_EINVAL, "acc_prover: null pointer");
    return curve == 0 ? acc_prover_dist<0>(instances, m, out_acc) : acc_prover_dist<1>(instances, m, out_acc);
HALO_API_END
}
int halo_dist_acc_verifier(int curve, const uint64_t* instances, size_t m, const uint64_t* acc) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(acc, HALO_EINVAL, "acc_verifier: null pointer");
    HALO_CHECK(dist_on(), HALO_ENOINIT, "halo_dist_init has not been called");
    return curve == 0 ? acc_verifier<0>(instances, m, acc, true) : acc_verifier<1>(instances, m, acc, true);
HALO_API_END
}
int halo_dist_acc_decider(int curve, const uint64_t* acc) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(acc, HALO_EINVAL, "acc_decider: null pointer");
    return curve == 0 ? pcdl_check_dist<0>(acc) : pcdl_check_dist<1>(acc);
HALO_API_END
}
int halo_hpoly_commit(int curve, const uint64_t* xis, size_t lg_n, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT(); NEED_CURVE();
    HALO_CHECK(xis && out_xyz && lg_n <= 30, HALO_EINVAL, "hpoly_commit: bad argument");
    HALO_CHECK(rt().pp[curve].n >= ((size_t)1 << lg_n), HALO_EPARAMS, "hpoly_commit: resident generators missing or too few");
    return curve == 0 ? hpoly_commit_words<0>(xis, (unsigned)lg_n, out_xyz) : hpoly_commit_words<1>(xis, (unsigned)lg_n, out_xyz);
HALO_API_END
}
int halo_acc_hpoly(int field, const uint64_t* xis, const uint64_t* alphas, size_t m, size_t lg_n, uint64_t* out_coeffs) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && xis && alphas && out_coeffs && lg_n <= 30, HALO_EINVAL, "acc_hpoly: bad argument");
    return field == 0 ? hpoly_coeffs<0>(xis, alphas, m, (unsigned)lg_n, out_coeffs) : hpoly_coeffs<1>(xis, alphas, m, (unsigned)lg_n, out_coeffs);
HALO_API_END
}
int halo_hpoly_eval(int field, const uint64_t* xis, size_t lg_n, const uint64_t z[4], uint64_t out[4]) { HALO_API_BEGIN
    HALO_CHECK((field == 0 || field == 1) && xis && z && out, HALO_EINVAL, "hpoly_eval: bad argument");
    if (field == 0) hpoly_eval_words<0>(xis, (unsigned)lg_n, z, out); else hpoly_eval_words<1>(xis, (unsigned)lg_n, z, out);
    return HALO_OK;
HALO_API_END
}
}  // extern "C"

extern "C" {
int halo_scalar_dot(int field, const uint64_t* xs, const uint64_t* ys, size_t n, uint64_t out[4]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && out && ((xs && ys) || n == 0), HALO_EINVAL, "scalar_dot: bad argument");
    return field == 0 ? scalar_dot_host<0>(xs, ys, n, out) : scalar_dot_host<1>(xs, ys, n, out);
HALO_API_END
}
int halo_construct_powers(int field, const uint64_t z[4], size_t n, uint64_t* out) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && z && (out || n == 0), HALO_EINVAL, "construct_powers: bad argument");
    return field == 0 ? construct_powers_host<0>(z, n, out) : construct_powers_host<1>(z, n, out);
HALO_API_END
}
int halo_poly_mul(int field, const uint64_t* a, size_t la, const uint64_t* b, size_t lb, uint64_t* out) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK((field == 0 || field == 1) && ((a && b && out) || la == 0 || lb == 0), HALO_EINVAL, "poly_mul: bad argument");
    HALO_CHECK(la + lb <= ((size_t)1 << 30), HALO_EINVAL, "poly_mul: product too long");
    return field == 0 ? poly_mul_host<0>(a, la, b, lb, out) : poly_mul_host<1>(a, la, b, lb, out);
HALO_API_END
}
}  // extern "C"
