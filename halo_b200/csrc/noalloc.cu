// halo_scalar_dot_noalloc_dev: the scalar_dot reduction (crates/group/src/group.rs:43-45; ipa_dot1_kernel in ipa.cu) for
// callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// It has no counterpart in the reference.
//
// The existing entry point (halo_scalar_dot) takes its per-block partial sums from the library's grow-only workspace
// arena (cudaMalloc on first use or growth), copies them to pinned host memory (cudaMallocHost on first use) and
// finishes the sum on the host.  This variant keeps the partial sums in a FILE-SCOPE device array of fixed size and
// finishes the sum on the device, so that one call is two kernel launches and nothing else: no cudaMalloc, no
// cudaFree, no host staging, no synchronisation.
//
// Limits and sharing:
//   * the scratch array holds NOALLOC_DOT_SLOTS = 4096 partial sums of 32 bytes (128 KiB; every GPU that loads the
//     library holds it until the process exits); one block reduces NOALLOC_DOT_PER_BLOCK = 4096 elements into one slot,
//     so n <= 4096 * 4096 = 2^24 elements per call.  Larger n is rejected with HALO_EINVAL.
//   * every call shares the array and it keeps the previous call's contents.  Nothing in it has to start at zero:
//     the second kernel reads exactly the slots the first kernel of the SAME call has written (ceil(n / 4096) of them,
//     none for n = 0), never a stale one.
//   * calls that overlap on the GPU (two streams, two host threads with their own streams, concurrent graph launches)
//     race on the array.  Calls on one stream are ordered and safe.
//
// In host code the array's name is only a host-side stand-in for the symbol, not a device address; the address the
// kernels receive comes from cudaGetSymbolAddress, taken once per device with that device current.
#include "runtime.cuh"

namespace halo {

constexpr unsigned NOALLOC_DOT_SLOTS = 4096;          // partial sums in the scratch array
constexpr unsigned NOALLOC_DOT_PER_BLOCK = 4096;      // elements one block reduces: 256 threads x 16
constexpr size_t NOALLOC_DOT_MAX_N = (size_t)NOALLOC_DOT_SLOTS * NOALLOC_DOT_PER_BLOCK;
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_noalloc_dot_scratch[NOALLOC_DOT_SLOTS * 32];

namespace {

template <int F> __device__ __forceinline__ Fe<F> na_ld(const Fe<F>* p) {
    Fe<F> r;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w;
    r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    return r;
}
template <int F> __device__ __forceinline__ void na_st(Fe<F>* p, const Fe<F>& v) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(v.l[0], v.l[1], v.l[2], v.l[3]);
    q[1] = make_uint4(v.l[4], v.l[5], v.l[6], v.l[7]);
}
// sum of the 256 per-thread values of a block (shared-memory tree); valid in thread 0
template <int F> __device__ __forceinline__ Fe<F> na_block_sum(Fe<F> s, Fe<F>* sh) {
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int d = 128; d >= 1; d >>= 1) {
        if ((int)threadIdx.x < d) sh[threadIdx.x] = fe_add<F>(sh[threadIdx.x], sh[threadIdx.x + d]);
        __syncthreads();
    }
    return sh[0];
}

// block b: part[b] = sum_{j in [4096 b, min(n, 4096 (b + 1)))} xs[j] * ys[j]; consecutive threads read consecutive
// elements (two 16-byte loads each: every warp request is a run of 1 KiB)
template <int F>
__global__ void __launch_bounds__(256) noalloc_dot_partials_kernel(const Fe<F>* __restrict__ xs, const Fe<F>* __restrict__ ys,
                                                                   size_t n, Fe<F>* __restrict__ part) {
    __shared__ Fe<F> sh[256];
    const size_t lo = (size_t)blockIdx.x * NOALLOC_DOT_PER_BLOCK;
    const size_t hi = lo + NOALLOC_DOT_PER_BLOCK < n ? lo + NOALLOC_DOT_PER_BLOCK : n;
    Fe<F> s = fe_zero<F>();
    for (size_t j = lo + threadIdx.x; j < hi; j += 256) s = fe_add<F>(s, fe_mul<F>(na_ld(xs + j), na_ld(ys + j)));
    s = na_block_sum<F>(s, sh);
    if (threadIdx.x == 0) na_st(part + blockIdx.x, s);
}
// one block: out = sum of the first nblk partial sums (zero for nblk = 0)
template <int F>
__global__ void __launch_bounds__(256) noalloc_dot_final_kernel(const Fe<F>* __restrict__ part, unsigned nblk, Fe<F>* __restrict__ out) {
    __shared__ Fe<F> sh[256];
    Fe<F> s = fe_zero<F>();
    for (unsigned j = threadIdx.x; j < nblk; j += 256) s = fe_add<F>(s, na_ld(part + j));
    s = na_block_sum<F>(s, sh);
    if (threadIdx.x == 0) na_st(out, s);
}

// device address of the scratch array, per device; looked up once with that device current
void* g_scratch_addr[NOALLOC_MAX_DEVICES] = {};

int noalloc_scratch_for_current_device(void** out) {
    int dev = -1;
    HALO_CUDA(cudaGetDevice(&dev));
    HALO_CHECK(dev >= 0 && dev < NOALLOC_MAX_DEVICES, HALO_EINVAL, "noalloc: device index out of range");
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        HALO_CUDA(cudaGetSymbolAddress(&p, g_noalloc_dot_scratch));
        HALO_CHECK(p != nullptr, HALO_ECUDA, "noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return HALO_OK;
}

}  // namespace
}  // namespace halo

using namespace halo;

extern "C" int halo_scalar_dot_noalloc_dev(int field, const void* d_xs, const void* d_ys, size_t n, void* d_out) {
    try {
        std::lock_guard<std::recursive_mutex> lock(rt().mu);
        HALO_CHECK(rt().inited, HALO_ENOINIT, "halo_init has not been called");
        HALO_CHECK(field == 0 || field == 1, HALO_EINVAL, "scalar_dot_noalloc: field must be 0 or 1");
        HALO_CHECK(d_out && ((d_xs && d_ys) || n == 0), HALO_EINVAL, "scalar_dot_noalloc: null pointer");
        HALO_CHECK(n <= NOALLOC_DOT_MAX_N, HALO_EINVAL,
                   "scalar_dot_noalloc: n exceeds the fixed scratch (at most 2^24 elements per call; use halo_scalar_dot)");
        Runtime& R = rt();
        HALO_CUDA(cudaSetDevice(R.device));                  // the device this library instance is bound to
        void* scratch = nullptr;
        if (int e = noalloc_scratch_for_current_device(&scratch)) return e;
        const unsigned nblk = (unsigned)((n + NOALLOC_DOT_PER_BLOCK - 1) / NOALLOC_DOT_PER_BLOCK);
        cudaStream_t st = R.stream;
        if (field == 0) {
            if (nblk) noalloc_dot_partials_kernel<0><<<nblk, 256, 0, st>>>((const Fe<0>*)d_xs, (const Fe<0>*)d_ys, n, (Fe<0>*)scratch);
            noalloc_dot_final_kernel<0><<<1, 256, 0, st>>>((const Fe<0>*)scratch, nblk, (Fe<0>*)d_out);
        } else {
            if (nblk) noalloc_dot_partials_kernel<1><<<nblk, 256, 0, st>>>((const Fe<1>*)d_xs, (const Fe<1>*)d_ys, n, (Fe<1>*)scratch);
            noalloc_dot_final_kernel<1><<<1, 256, 0, st>>>((const Fe<1>*)scratch, nblk, (Fe<1>*)d_out);
        }
        R.launches += nblk ? 2 : 1;
        HALO_CUDA(cudaGetLastError());
        return HALO_OK;
    } catch (...) {
        halo::last_err() = "scalar_dot_noalloc: unexpected exception";
        return HALO_ECUDA;
    }
}
