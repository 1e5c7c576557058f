// Poseidon on the device: many INDEPENDENT sponges at once, one thread each.
//
// Restates crates/poseidon/src/inner_sponge.rs:3-115 (x^7 s-box, 55 full rounds of sbox -> MDS -> + round constant,
// width 3 = rate 2 + capacity 1; absorb adds into the rate cells and permutes when both are full; squeeze permutes when
// absorbing or when both cells were handed out) for the device, with the constants of
// crates/group/src/poseidon_consts.rs (poseidon_consts.inc, data only).  SURVEY.md section 8(f) #4.
//
// Where it pays and where it does not: ONE sponge is a serial chain of ~1150 dependent
// multiplications per permutation -- 70 us on a GPU thread against 17 us on a host core -- so the Fiat-Shamir transcript
// of an opening stays on the host.  Independent sponges are a different matter: n of them cost one permutation's latency
// on the device (2^16 hashes of two elements: ~0.2 ms against ~70 ms on sixteen host threads), which is what a batch of
// transcripts or a Merkle / hash-to-field layer over a column needs.
#include "runtime.cuh"
#include <cstring>

namespace halo {

namespace poseidon_dev {
#include "poseidon_consts.inc"
__constant__ unsigned long long c_mds[2][9][4];
__constant__ unsigned long long c_rc[2][165][4];
}  // namespace poseidon_dev

template <int F> __device__ __forceinline__ Fe<F> c_fe(const unsigned long long* w) {
    Fe<F> r;
#pragma unroll
    for (int i = 0; i < 4; ++i) { r.l[2 * i] = (u32)w[i]; r.l[2 * i + 1] = (u32)(w[i] >> 32); }
    return r;
}

template <int F> __device__ void poseidon_permute_dev(Fe<F> (&s)[3]) {
    for (int r = 0; r < 55; ++r) {
        Fe<F> t[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const Fe<F> x2 = fe_sqr<F>(s[i]), x3 = fe_mul<F>(s[i], x2), x4 = fe_sqr<F>(x2);
            t[i] = fe_mul<F>(x3, x4);
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            Fe<F> acc = c_fe<F>(poseidon_dev::c_rc[F][3 * r + i]);
#pragma unroll
            for (int j = 0; j < 3; ++j) acc = fe_add<F>(acc, fe_mul<F>(t[j], c_fe<F>(poseidon_dev::c_mds[F][3 * i + j])));
            s[i] = acc;
        }
    }
}

// sponge k absorbs inputs[k * len .. (k + 1) * len) and squeezes one element
template <int F>
__global__ void __launch_bounds__(128) poseidon_hash_kernel(const Fe<F>* __restrict__ inputs, size_t len, size_t count,
                                                            Fe<F>* __restrict__ out) {
    const size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (k >= count) return;
    Fe<F> s[3] = {fe_zero<F>(), fe_zero<F>(), fe_zero<F>()};
    int filled = 0;
    const uint4* in = reinterpret_cast<const uint4*>(inputs + k * len);
    for (size_t i = 0; i < len; ++i) {
        if (filled == 2) { poseidon_permute_dev<F>(s); filled = 0; }
        const uint4 a = in[2 * i], b = in[2 * i + 1];
        Fe<F> x;
        x.l[0] = a.x; x.l[1] = a.y; x.l[2] = a.z; x.l[3] = a.w; x.l[4] = b.x; x.l[5] = b.y; x.l[6] = b.z; x.l[7] = b.w;
        if (filled == 0) s[0] = fe_add<F>(s[0], x); else s[1] = fe_add<F>(s[1], x);
        ++filled;
    }
    poseidon_permute_dev<F>(s);                            // squeeze after absorbing always permutes first
    uint4* o = reinterpret_cast<uint4*>(out + k);
    o[0] = make_uint4(s[0].l[0], s[0].l[1], s[0].l[2], s[0].l[3]);
    o[1] = make_uint4(s[0].l[4], s[0].l[5], s[0].l[6], s[0].l[7]);
}

static int poseidon_constants() {
    static int loaded_device = -1;
    Runtime& R = rt();
    if (loaded_device == R.device) return HALO_OK;
    HALO_CUDA(cudaMemcpyToSymbol(poseidon_dev::c_mds, poseidon_dev::POSEIDON_MDS, sizeof(poseidon_dev::POSEIDON_MDS)));
    HALO_CUDA(cudaMemcpyToSymbol(poseidon_dev::c_rc, poseidon_dev::POSEIDON_RC, sizeof(poseidon_dev::POSEIDON_RC)));
    loaded_device = R.device;
    return HALO_OK;
}

// count independent sponges of len elements each, device buffers
int poseidon_hash_batch_device(int field, const void* d_inputs, size_t len, size_t count, void* d_out) {
    Runtime& R = rt();
    if (count == 0) return HALO_OK;
    if (int e = poseidon_constants()) return e;
    const unsigned grid = (unsigned)((count + 127) / 128);
    if (field == 0) poseidon_hash_kernel<0><<<grid, 128, 0, R.stream>>>((const Fe<0>*)d_inputs, len, count, (Fe<0>*)d_out);
    else poseidon_hash_kernel<1><<<grid, 128, 0, R.stream>>>((const Fe<1>*)d_inputs, len, count, (Fe<1>*)d_out);
    R.launches += 1;
    HALO_CUDA(cudaGetLastError());
    return HALO_OK;
}

}  // namespace halo

using namespace halo;
#define LOCK() std::lock_guard<std::recursive_mutex> lock__(rt().mu)
#define NEED_INIT() HALO_CHECK(rt().inited, HALO_ENOINIT, "halo_init has not been called")
#define HALO_API_BEGIN try {
#define HALO_API_END } catch (const std::bad_alloc&) { halo::last_err() = "out of host memory"; return HALO_EINVAL; } \
    catch (const std::exception& e__) { halo::last_err() = std::string("exception: ") + e__.what(); return HALO_ECUDA; } \
    catch (...) { halo::last_err() = "unknown exception"; return HALO_ECUDA; }

extern "C" {

int halo_poseidon_hash_batch(int field, const uint64_t* inputs, size_t len, size_t count, uint64_t* out) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(field == 0 || field == 1, HALO_EINVAL, "field must be 0 (Fp) or 1 (Fq)");
    HALO_CHECK(out || count == 0, HALO_EINVAL, "poseidon: null output");
    HALO_CHECK(inputs || len == 0 || count == 0, HALO_EINVAL, "poseidon: null input");
    HALO_CHECK(len <= ((size_t)1 << 24) && count <= ((size_t)1 << 28), HALO_EINVAL, "poseidon: batch too large");
    if (count == 0) return HALO_OK;
    Runtime& R = rt();
    void *d_in = nullptr, *d_out = nullptr;
    if (int e = ws_get(WS_TMP0, std::max<size_t>(32, count * len * 32), &d_in)) return e;
    if (int e = ws_get(WS_TMP1, count * 32, &d_out)) return e;
    if (len) HALO_CUDA(cudaMemcpyAsync(d_in, inputs, count * len * 32, cudaMemcpyHostToDevice, R.stream));
    if (int e = poseidon_hash_batch_device(field, d_in, len, count, d_out)) return e;
    HALO_CUDA(cudaMemcpyAsync(out, d_out, count * 32, cudaMemcpyDeviceToHost, R.stream));
    HALO_CUDA(cudaStreamSynchronize(R.stream));
    return HALO_OK;
HALO_API_END
}

int halo_poseidon_hash_batch_dev(int field, const void* d_inputs, size_t len, size_t count, void* d_out) { HALO_API_BEGIN
    LOCK(); NEED_INIT();
    HALO_CHECK(field == 0 || field == 1, HALO_EINVAL, "field must be 0 (Fp) or 1 (Fq)");
    HALO_CHECK((d_out && (d_inputs || len == 0)) || count == 0, HALO_EINVAL, "poseidon: null pointer");
    return poseidon_hash_batch_device(field, d_inputs, len, count, d_out);
HALO_API_END
}

}  // extern "C"
