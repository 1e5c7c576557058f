// Process-wide runtime state of libhalo_b200: device binding, stream, resident generator sets (the device-side
// analogue of the reference's per-curve OnceLock<PublicParams>, crates/group/src/wrappers.rs:28-29) and a
// grow-only workspace arena so that steady-state calls do not allocate.
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <string>
#include <vector>
#include <thread>
#include <functional>
#include <cstdio>
#include "curve.cuh"
#include "../../include/halo_b200.h"

namespace halo {

struct DeviceBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct ResidentParams {
    void* d_gs = nullptr;      // n affine points, 64 B each
    size_t n = 0;
    uint64_t s_xy[8], h_xy[8]; // host copies (affine, Montgomery)
    void* d_fold_tab = nullptr;   // IPA fixed-base fold table: 2^k * Gs[n/2 + j], k-major (ipa.cu), for openings of size fold_tab_n
    size_t fold_tab_n = 0;
    void* d_fixed_tab = nullptr;      // commit path: rows 2^(c w) * Gs[i] (ipa.cu ensure_fixed_table, msm.cu msm_fixed)
    size_t fixed_tab_n = 0;
    int fixed_tab_c = 0;
    // the same table over this rank's generators of a sharded opening (ipa.cu): slot 0 for the whole communicator, slot 1
    // for its half (the one-proof prover uses both in every proof)
    void* d_fold_tab_loc[2] = {nullptr, nullptr};
    size_t fold_tab_loc_n[2] = {0, 0};
    int fold_tab_loc_world[2] = {0, 0}, fold_tab_loc_rank[2] = {0, 0};
};

struct Runtime {
    bool inited = false;
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    cudaStream_t copy_stream = nullptr;   // second stream: the base-point upload of halo_msm overlaps the scalar pipeline
    cudaEvent_t copy_done = nullptr, copy_go = nullptr;
    bool wait_copy = false;               // the next MSM must wait for copy_done before touching its bases
    // pageable caller buffers: the base points are staged to the device by a helper thread while the calling thread
    // already enqueues the scalar half of the pipeline; joined (all chunks ISSUED, copy_done recorded) right before the
    // first kernel that reads the bases
    std::thread upload_thread;
    int upload_rc = 0;
    // chunked host MSM (msm_host_chunked): one event per uploaded piece, and the action the pipeline runs right before
    // its first kernel that reads the bases
    cudaEvent_t chunk_ev[32] = {};
    std::function<int(cudaStream_t)> bases_hook;
    ResidentParams pp[2];
    std::vector<DeviceBuf> arena;   // indexed slots
    void* pinned = nullptr;         // pinned host staging
    size_t pinned_cap = 0;
    // staged upload of large pageable buffers: UP_THREADS workers, two pinned chunks and one stream each
    static constexpr int UP_THREADS = 16, UP_SLOTS = 2;
    static constexpr size_t UP_CHUNK = (size_t)2 << 20;
    void* up_buf[UP_THREADS][UP_SLOTS] = {};
    cudaEvent_t up_ev[UP_THREADS][UP_SLOTS] = {};
    cudaStream_t up_stream[UP_THREADS] = {};
    int msm_window_override = 0;
    int tune[16] = {};              // experiment knobs (halo_set_tuning)
    // per-stage CUDA-event timing of the last MSM (halo_profile_*), and a count of kernels launched
    bool prof_on = false;
    static const int PROF_MAX = 24;
    cudaEvent_t prof_ev[PROF_MAX + 1] = {};
    const char* prof_name[PROF_MAX] = {};
    int prof_n = 0;
    unsigned long long launches = 0;
    void* ntt_tw[2] = {nullptr, nullptr};   // twiddle table per field: omega_max^k, k < 2^(logmax-1)
    int ntt_logmax[2] = {0, 0};
    std::recursive_mutex mu;
    bool is_lane = false;           // a second execution context (lane_get): shares the resident tables, never builds them
};

// The runtime of the calling thread: the process-wide one, or -- inside a LaneScope -- the second execution context.
Runtime& rt();
// Second execution context on the same device: its own stream, workspace arena and pinned buffer, with the resident
// generator sets, fold tables, twiddle tables and tuning knobs of the main runtime copied in (pointers, not data) at
// every call.  A helper thread that enters it (LaneScope) can run one library pipeline -- an IPA opening -- beside
// the calling thread's: the two independent Instance::open calls of a PLONK proof (protocol.rs:295-296) are each a
// chain of latency-bound rounds that leave most of the GPU idle.  Returns nullptr when it cannot be created.
Runtime* lane_get();
void lane_finish(Runtime* lane);    // adds the lane's launch count to the main runtime
struct LaneScope {
    explicit LaneScope(Runtime* r);
    ~LaneScope();
};
// last error message of the calling thread (halo_last_error)
inline std::string& last_err() { static thread_local std::string e; return e; }

#define HALO_CUDA(call)                                                                        \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            char buf__[512];                                                                   \
            snprintf(buf__, sizeof buf__, "%s:%d: %s -> %s", __FILE__, __LINE__, #call,        \
                     cudaGetErrorString(e__));                                                 \
            halo::last_err() = buf__;                                                            \
            return HALO_ECUDA;                                                                 \
        }                                                                                      \
    } while (0)

#define HALO_CHECK(cond, code, msg)                 \
    do {                                            \
        if (!(cond)) {                              \
            halo::last_err() = (msg);                 \
            return (code);                          \
        }                                           \
    } while (0)

// Worker threads that are always joined: thread creation failing half way (std::system_error) or an exception in the
// spawning thread must not destroy a joinable std::thread, which would terminate the host process.
struct Workers {
    std::vector<std::thread> th;
    template <class F> void spawn(unsigned n, F f) {
        for (unsigned i = 0; i < n; ++i) {
            try { th.emplace_back(f); } catch (...) { break; }
        }
    }
    template <class F> void spawn_indexed(unsigned first, unsigned last, F f) {   // f(t) for t in [first, last)
        for (unsigned t = first; t < last; ++t) {
            try { th.emplace_back(f, t); } catch (...) { for (; t < last; ++t) f(t); break; }
        }
    }
    size_t size() const { return th.size(); }
    void join() { for (auto& t : th) if (t.joinable()) t.join(); th.clear(); }
    ~Workers() { join(); }
};


// stage timing: prof_begin() at the start of a pipeline, prof_mark("name") after each stage's launches
int prof_begin();
int prof_mark(const char* name);

// workspace slot: returns a device pointer with at least `bytes` capacity (contents undefined)
int ws_get(int slot, size_t bytes, void** out);
int pinned_get(size_t bytes, void** out);
// host -> device copy of a (pageable) caller buffer, ordered after everything already queued on the library stream
// and complete on that stream when it returns: worker threads stage chunks through pinned memory so the CPU copy
// and the DMA overlap and several cores feed the link
int upload_staged(void* d_dst, const void* h_src, size_t bytes, cudaStream_t order_stream = nullptr);
// true when p is ordinary (pageable, unregistered) host memory, i.e. what a Rust Vec or a numpy array is
bool host_ptr_is_pageable(const void* p);
// host scalars (and optionally bases) of an MSM -> the WS_SCALARS / WS_BASES workspaces, bases overlapping the scalar
// half of the pipeline (see wait_copy / upload_thread); msm_upload_finish joins and clears after the MSM returned
int msm_upload_inputs(const uint64_t* bases_xy, size_t base_off, const uint64_t* scalars, size_t k, void** d_s, void** d_b);
void msm_upload_finish();
// called by the MSM pipelines right before the first kernel that reads the bases
int msm_wait_bases(cudaStream_t st);

enum WsSlot {
    WS_SCALARS = 0, WS_BASES, WS_COUNTS, WS_OFFSETS, WS_CURSOR, WS_IDX, WS_ORDER, WS_SIZEHIST, WS_BUCKETS,
    WS_SEGSUM, WS_SEGWGT, WS_WINSUM, WS_TMP0, WS_TMP1, WS_TMP2, WS_TMP3, WS_IPA_G, WS_IPA_C, WS_IPA_Z, WS_IPA_T0,
    WS_IPA_T1, WS_IPA_S, WS_IPA_W, WS_NTT0, WS_NTT1, WS_PLONK_COLS, WS_PLONK_TMP, WS_PLONK_EV, WS_PLONK_F0,
    WS_PLONK_F1, WS_PLONK_MISC, WS_NTT_TAB, WS_FOLD_TMP, WS_FOLD_STATE, WS_BA_W, WS_BA_P, WS_PLONK_PACK, WS_IPA_GL, WS_IPA_CL, WS_COUNT
};

}  // namespace halo
