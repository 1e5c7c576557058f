// Multi-GPU layer of libhalo_b200: one process per GPU, an NCCL communicator owned by the library.
//
// the sums on this path shard by point range (MSM), by instance (AS_DL) and by polynomial
// (NTT batches).  The exchange steps are tiny: a point-range-sharded MSM leaves nwin window sums (2 KB) per rank, which are
// all-gathered straight out of the device buffer the reduction wrote them to (no host bounce in front of the
// collective), copied to the host once and combined there (N * nwin additions and ONE Horner: a GPU thread needs
// ~1 ms for the 255 dependent doublings, a host core 0.05 ms).  Elliptic-curve addition and modular addition are not
// NCCL reduction operators, so every collective here is an all-gather followed by our own combine.
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the process already has, e.g. torch's, or the system
// one), so the single-GPU path has no NCCL dependency at all.
#include "runtime.cuh"
#include "host.cuh"
#include <nccl.h>
#include <dlfcn.h>
#include <algorithm>
#include <cstring>

namespace halo {

template <int C> int msm_windows_device(const void* d_bases, const void* d_scalars, size_t k, size_t nbatch, ptrdiff_t base_stride,
                                        void** d_winsum, int* c_out, int* nwin_out);
template <int C> void msm_horner_host(const void* win, int c, int nwin, uint64_t out_xyz[12]);
template <int C> int msm_host_chunked_windows(const uint64_t* h_bases, const uint64_t* h_scalars, size_t k, void** d_winsum, int* c_out, int* nwin_out);

namespace {
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t*, void*) = nullptr;      // optional (NCCL >= 2.18)
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
struct DistState {
    bool on = false;
    int rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    NcclApi api;
    void* d_send = nullptr; void* d_recv = nullptr; size_t send_cap = 0, recv_cap = 0;
    void* h_recv = nullptr; size_t h_cap = 0;
    // the two HALVES of the communicator (ranks [0, world/2) and [world/2, world)), created on first use: inside a
    // half scope (dist_half_enter / _leave) rank, world and every collective of this file refer to the caller's half
    ncclComm_t full_comm = nullptr, half_comm = nullptr;
    int full_rank = 0, full_world = 1;
    bool in_half = false;
};
DistState& ds() { static DistState s; return s; }

int nccl_load() {
    NcclApi& a = ds().api;
    if (a.handle) return HALO_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    HALO_CHECK(h != nullptr, HALO_ECUDA, "dist: libnccl.so.2 not found");
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
    a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    a.Broadcast = (decltype(a.Broadcast))dlsym(h, "ncclBroadcast");
    a.CommSplit = (decltype(a.CommSplit))dlsym(h, "ncclCommSplit");
    HALO_CHECK(a.GetUniqueId && a.CommInitRank && a.CommDestroy && a.AllGather && a.Broadcast && a.GetErrorString, HALO_ECUDA,
               "dist: libnccl lacks an expected symbol");
    a.handle = h;
    return HALO_OK;
}
#define HALO_NCCL(call)                                                                                   \
    do {                                                                                                  \
        ncclResult_t r__ = (call);                                                                        \
        if (r__ != ncclSuccess) {                                                                         \
            halo::last_err() = std::string("nccl: ") + #call + " -> " + ds().api.GetErrorString(r__);     \
            return HALO_ECUDA;                                                                            \
        }                                                                                                 \
    } while (0)

int grow(void** p, size_t* cap, size_t bytes, bool host) {
    if (*cap >= bytes) return HALO_OK;
    HALO_CUDA(cudaStreamSynchronize(rt().stream));
    if (*p) { if (host) cudaFreeHost(*p); else cudaFree(*p); *p = nullptr; *cap = 0; }
    const size_t want = bytes + bytes / 4 + 4096;
    if (host) HALO_CUDA(cudaMallocHost(p, want)); else HALO_CUDA(cudaMalloc(p, want));
    *cap = want;
    return HALO_OK;
}
}  // namespace

bool dist_on() { return ds().on; }
int dist_rank() { return ds().rank; }
int dist_world() { return ds().world; }

// Half scope: two independent collectives (the two openings of one proof) run side by side on the two halves of the box.
bool dist_half_available() {
    const DistState& D = ds();
    return D.on && !D.in_half && D.world >= 4 && D.world % 2 == 0 && D.api.CommSplit != nullptr;
}
int dist_half_enter() {                                    // collective over the FULL communicator on first use
    DistState& D = ds();
    HALO_CHECK(dist_half_available(), HALO_EINVAL, "dist: no half communicator (needs an even world >= 4 and ncclCommSplit)");
    if (!D.half_comm) {
        HALO_CUDA(cudaStreamSynchronize(rt().stream));
        HALO_NCCL(D.api.CommSplit(D.comm, D.rank / (D.world / 2), D.rank, &D.half_comm, nullptr));
    }
    D.full_comm = D.comm; D.full_rank = D.rank; D.full_world = D.world;
    D.comm = D.half_comm; D.world = D.full_world / 2; D.rank = D.full_rank % D.world;
    D.in_half = true;
    return HALO_OK;
}
void dist_half_leave() {
    DistState& D = ds();
    if (!D.in_half) return;
    D.comm = D.full_comm; D.rank = D.full_rank; D.world = D.full_world;
    D.in_half = false;
}
bool dist_in_half() { return ds().in_half; }

// all-gather of `bytes` per rank between device buffers, on the library stream
int dist_allgather_dev(const void* d_send, void* d_recv, size_t bytes) {
    DistState& D = ds();
    HALO_CHECK(D.on, HALO_ENOINIT, "halo_dist_init has not been called");
    HALO_NCCL(D.api.AllGather(d_send, d_recv, bytes, ncclUint8, D.comm, rt().stream));
    return HALO_OK;
}
// all-gather of host data (small control payloads: challenge vectors, points): bounced through device staging
int dist_allgather_host(const void* h_send, void* h_recv_all, size_t bytes) {
    DistState& D = ds();
    HALO_CHECK(D.on, HALO_ENOINIT, "halo_dist_init has not been called");
    cudaStream_t st = rt().stream;
    if (int e = grow(&D.d_send, &D.send_cap, bytes, false)) return e;
    if (int e = grow(&D.d_recv, &D.recv_cap, bytes * D.world, false)) return e;
    HALO_CUDA(cudaMemcpyAsync(D.d_send, h_send, bytes, cudaMemcpyHostToDevice, st));
    HALO_NCCL(D.api.AllGather(D.d_send, D.d_recv, bytes, ncclUint8, D.comm, st));
    HALO_CUDA(cudaMemcpyAsync(h_recv_all, D.d_recv, bytes * D.world, cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    return HALO_OK;
}

// broadcast of a device buffer from `root`, in place, on the library stream
int dist_bcast_dev(void* d_buf, size_t bytes, int root) {
    DistState& D = ds();
    HALO_CHECK(D.on, HALO_ENOINIT, "halo_dist_init has not been called");
    if (bytes == 0) return HALO_OK;
    HALO_NCCL(D.api.Broadcast(d_buf, d_buf, bytes, ncclUint8, root, D.comm, rt().stream));
    return HALO_OK;
}
// broadcast of host data (small control payloads) from `root`, bounced through device staging; h_buf is read on the
// root and written on the other ranks
int dist_bcast_host(void* h_buf, size_t bytes, int root) {
    DistState& D = ds();
    HALO_CHECK(D.on, HALO_ENOINIT, "halo_dist_init has not been called");
    if (bytes == 0) return HALO_OK;
    cudaStream_t st = rt().stream;
    if (int e = grow(&D.d_send, &D.send_cap, bytes, false)) return e;
    if (D.rank == root) HALO_CUDA(cudaMemcpyAsync(D.d_send, h_buf, bytes, cudaMemcpyHostToDevice, st));
    HALO_NCCL(D.api.Broadcast(D.d_send, D.d_send, bytes, ncclUint8, root, D.comm, st));
    if (D.rank != root) HALO_CUDA(cudaMemcpyAsync(h_buf, D.d_send, bytes, cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    return HALO_OK;
}

// this rank's share [lo, hi) of `total` items, contiguous and balanced
void dist_range(size_t total, size_t* lo, size_t* hi) {
    const size_t w = (size_t)ds().world, r = (size_t)ds().rank;
    *lo = total * r / w; *hi = total * (r + 1) / w;
}

// ---------------------------------------------------------------------------------------------------------
// point-range-sharded MSM.  Payload of one rank: a 128-byte header and, per MSM of the batch, 32 window slots.
// ---------------------------------------------------------------------------------------------------------
struct MsmHdr { u32 c, nwin, nbatch, pad[29]; };
static_assert(sizeof(MsmHdr) == 128, "header is one XYZZ slot");
constexpr int DIST_SLOTS = 32;           // >= windows of the narrowest width in use (c = 8: 32 windows)

template <int C>
__global__ void dist_pack_kernel(const XYZZ<C>* __restrict__ winsum, u32 c, u32 nwin, u32 nbatch, unsigned char* __restrict__ out) {
    if (blockIdx.x == 0 && threadIdx.x < 32) {
        u32* h = reinterpret_cast<u32*>(out);
        h[threadIdx.x] = threadIdx.x == 0 ? c : threadIdx.x == 1 ? nwin : threadIdx.x == 2 ? nbatch : 0u;
    }
    // one u32 per thread: slot (j, w) word t  <-  winsum[j * nwin + w] word t
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    const u32 words = nbatch * nwin * 32;
    if (t >= words) return;
    const u32 item = t / 32, word = t % 32, j = item / nwin, w = item % nwin;
    reinterpret_cast<u32*>(out + 128)[(j * DIST_SLOTS + w) * 32 + word] = reinterpret_cast<const u32*>(winsum)[t];
}

// base_stride != 0: MSM j of the batch reads bases + j * base_stride (the IPA's L / R pair over the two halves)
// h_bases != nullptr: this rank's slice comes from HOST buffers (one MSM) and is uploaded pipelined with the
// accumulation by point range (msm_host_chunked_windows) -- d_bases / d_scalars are then the host pointers
template <int C>
int dist_msm_device(const void* d_bases, const void* d_scalars, size_t k_local, size_t nbatch, uint64_t* out_xyz, ptrdiff_t base_stride,
                    bool from_host) {
    DistState& D = ds();
    Runtime& R = rt();
    HALO_CHECK(D.on, HALO_ENOINIT, "halo_dist_init has not been called");
    HALO_CHECK(nbatch >= 1 && nbatch <= 64, HALO_EINVAL, "dist msm: batch must be in [1, 64]");
    cudaStream_t st = R.stream;
    const size_t bytes = 128 + nbatch * DIST_SLOTS * sizeof(XYZZ<C>);
    if (int e = grow(&D.d_send, &D.send_cap, bytes, false)) return e;
    if (int e = grow(&D.d_recv, &D.recv_cap, bytes * D.world, false)) return e;
    if (int e = grow(&D.h_recv, &D.h_cap, bytes * D.world, true)) return e;
    if (k_local == 0) {
        HALO_CUDA(cudaMemsetAsync(D.d_send, 0, 128, st));                      // nwin = 0: nothing to add
    } else {
        void* d_win = nullptr; int c = 0, nwin = 0;
        if (from_host) {
            if (int e = msm_host_chunked_windows<C>((const uint64_t*)d_bases, (const uint64_t*)d_scalars, k_local, &d_win, &c, &nwin)) return e;
        } else
        if (int e = msm_windows_device<C>(d_bases, d_scalars, k_local, nbatch, base_stride, &d_win, &c, &nwin)) return e;
        HALO_CHECK(nwin <= DIST_SLOTS, HALO_EINVAL, "dist msm: window count above the payload slots");
        const u32 words = (u32)(nbatch * nwin * 32);
        dist_pack_kernel<C><<<(words + 255) / 256, 256, 0, st>>>((const XYZZ<C>*)d_win, (u32)c, (u32)nwin, (u32)nbatch, (unsigned char*)D.d_send);
        R.launches += 1;
        HALO_CUDA(cudaGetLastError());
    }
    HALO_NCCL(D.api.AllGather(D.d_send, D.d_recv, bytes, ncclUint8, D.comm, st));
    HALO_CUDA(cudaMemcpyAsync(D.h_recv, D.d_recv, bytes * D.world, cudaMemcpyDeviceToHost, st));
    HALO_CUDA(cudaStreamSynchronize(st));
    // combine: ranks with the same window width are added window by window, then one Horner per width in use
    const unsigned char* hr = (const unsigned char*)D.h_recv;
    for (size_t j = 0; j < nbatch; ++j) {
        XYZZ<C> total = xyzz_identity<C>();
        bool used[17] = {};
        for (int r0 = 0; r0 < D.world; ++r0) {
            const MsmHdr* h0 = (const MsmHdr*)(hr + (size_t)r0 * bytes);
            if (h0->nwin == 0 || h0->c > 16 || used[h0->c]) continue;
            HALO_CHECK(h0->nwin <= (u32)DIST_SLOTS && h0->nbatch == nbatch, HALO_ECUDA, "dist msm: ranks disagree on the batch shape");
            used[h0->c] = true;
            XYZZ<C> win[DIST_SLOTS];
            for (u32 w = 0; w < h0->nwin; ++w) win[w] = xyzz_identity<C>();
            for (int r = r0; r < D.world; ++r) {
                const MsmHdr* h = (const MsmHdr*)(hr + (size_t)r * bytes);
                if (h->nwin == 0 || h->c != h0->c) continue;
                const XYZZ<C>* src = (const XYZZ<C>*)((const unsigned char*)h + 128) + j * DIST_SLOTS;
                for (u32 w = 0; w < h->nwin; ++w) xyzz_add<C>(win[w], src[w]);
            }
            uint64_t part[12];
            msm_horner_host<C>(win, (int)h0->c, (int)h0->nwin, part);
            xyzz_add<C>(total, Host<C>::from_words(part));
        }
        Host<C>::to_words(total, out_xyz + 12 * j);
    }
    return HALO_OK;
}
template int dist_msm_device<0>(const void*, const void*, size_t, size_t, uint64_t*, ptrdiff_t, bool);
template int dist_msm_device<1>(const void*, const void*, size_t, size_t, uint64_t*, ptrdiff_t, bool);

}  // namespace halo

using namespace halo;
#define LOCK() std::lock_guard<std::recursive_mutex> lock__(rt().mu)
#define NEED_INIT() HALO_CHECK(rt().inited, HALO_ENOINIT, "halo_init has not been called")
#define HALO_API_BEGIN try {
#define HALO_API_END } catch (const std::bad_alloc&) { halo::last_err() = "out of host memory"; return HALO_EINVAL; } \
    catch (const std::exception& e__) { halo::last_err() = std::string("exception: ") + e__.what(); return HALO_ECUDA; } \
    catch (...) { halo::last_err() = "unknown exception"; return HALO_ECUDA; }

extern "C" {

int halo_dist_unique_id(uint8_t out[128]) { HALO_API_BEGIN
    LOCK();
    HALO_CHECK(out, HALO_EINVAL, "dist_unique_id: null pointer");
    if (int e = nccl_load()) return e;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    HALO_NCCL(ds().api.GetUniqueId(&id));
    memcpy(out, &id, 128);
    return HALO_OK;
HALO_API_END
}

int halo_dist_init(int rank, int world, const uint8_t id_bytes[128]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    DistState& D = ds();
    HALO_CHECK(id_bytes && world >= 1 && rank >= 0 && rank < world, HALO_EINVAL, "dist_init: bad rank / world / id");
    if (D.on) { HALO_CHECK(D.rank == rank && D.world == world, HALO_EINVAL, "dist_init: already initialised with another shape"); return HALO_OK; }
    if (int e = nccl_load()) return e;
    HALO_CUDA(cudaSetDevice(rt().device));
    ncclUniqueId id;
    memcpy(&id, id_bytes, 128);
    HALO_NCCL(D.api.CommInitRank(&D.comm, world, id, rank));
    D.rank = rank; D.world = world; D.on = true;
    return HALO_OK;
HALO_API_END
}

int halo_dist_shutdown(void) { HALO_API_BEGIN
    LOCK();
    DistState& D = ds();
    if (!D.on) return HALO_OK;
    if (rt().inited) cudaStreamSynchronize(rt().stream);
    dist_half_leave();
    if (D.half_comm) { D.api.CommDestroy(D.half_comm); D.half_comm = nullptr; }
    D.api.CommDestroy(D.comm);
    if (D.d_send) cudaFree(D.d_send);
    if (D.d_recv) cudaFree(D.d_recv);
    if (D.h_recv) cudaFreeHost(D.h_recv);
    D.d_send = D.d_recv = D.h_recv = nullptr; D.send_cap = D.recv_cap = D.h_cap = 0;
    D.comm = nullptr; D.on = false; D.rank = 0; D.world = 1;
    return HALO_OK;
HALO_API_END
}

int halo_dist_info(int* rank, int* world) {
    if (rank) *rank = ds().rank;
    if (world) *world = ds().world;
    return ds().on ? HALO_OK : HALO_ENOINIT;
}

int halo_dist_msm_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k_local, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (d_scalars || k_local == 0), HALO_EINVAL, "dist msm: null pointer");
    Runtime& R = rt();
    const char* bases = (const char*)d_bases_xy;
    if (!bases) {
        HALO_CHECK(R.pp[curve].n >= base_off + k_local, HALO_EPARAMS, "dist msm: resident generators missing or too few");
        bases = (const char*)R.pp[curve].d_gs;
    }
    bases += base_off * 64;
    return curve == 0 ? dist_msm_device<0>(bases, d_scalars, k_local, 1, out_xyz, 0, false) : dist_msm_device<1>(bases, d_scalars, k_local, 1, out_xyz, 0, false);
HALO_API_END
}

int halo_dist_msm(int curve, const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k_local, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (scalars || k_local == 0), HALO_EINVAL, "dist msm: null pointer");
    // a large slice over caller-supplied bases: upload pipelined with the accumulation by point range, as halo_msm does
    if (bases_xy && k_local >= ((size_t)1 << 19) && k_local < ((size_t)1 << 27) && rt().tune[9] != 1 && rt().tune[8] == 0)
        return curve == 0 ? dist_msm_device<0>(bases_xy + 8 * base_off, scalars, k_local, 1, out_xyz, 0, true)
                          : dist_msm_device<1>(bases_xy + 8 * base_off, scalars, k_local, 1, out_xyz, 0, true);
    void *d_s = nullptr, *d_b = nullptr;
    if (int e = msm_upload_inputs(bases_xy, base_off, scalars, k_local, &d_s, &d_b)) { msm_upload_finish(); return e; }
    if (d_b) base_off = 0;
    int rc = halo_dist_msm_dev(curve, d_b, base_off, d_s, k_local, out_xyz);
    msm_upload_finish();
    return rc;
HALO_API_END
}

int halo_dist_msm_global(int curve, const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k, uint64_t out_xyz[12]) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(ds().on, HALO_ENOINIT, "halo_dist_init has not been called");
    HALO_CHECK(out_xyz && (scalars || k == 0), HALO_EINVAL, "dist msm: null pointer");
    size_t lo, hi;
    dist_range(k, &lo, &hi);
    return halo_dist_msm(curve, bases_xy, base_off + lo, scalars ? scalars + 4 * lo : nullptr, hi - lo, out_xyz);
HALO_API_END
}

int halo_dist_msm_batch_dev(int curve, const void* d_bases_xy, size_t base_off, const void* d_scalars, size_t k_local, size_t b,
                            uint64_t* out_xyz) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(curve == 0 || curve == 1, HALO_EINVAL, "bad curve");
    HALO_CHECK(out_xyz && (d_scalars || k_local == 0) && b >= 1, HALO_EINVAL, "dist msm batch: bad argument");
    Runtime& R = rt();
    const char* bases = (const char*)d_bases_xy;
    if (!bases) {
        HALO_CHECK(R.pp[curve].n >= base_off + k_local, HALO_EPARAMS, "dist msm batch: resident generators missing or too few");
        bases = (const char*)R.pp[curve].d_gs;
    }
    bases += base_off * 64;
    return curve == 0 ? dist_msm_device<0>(bases, d_scalars, k_local, b, out_xyz, 0, false) : dist_msm_device<1>(bases, d_scalars, k_local, b, out_xyz, 0, false);
HALO_API_END
}

}  // extern "C"
