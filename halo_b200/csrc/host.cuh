// Host-side group and scalar helpers shared by the protocol layers (O(lg n) work only), and the entry points the
// translation units call across each other.
#pragma once
#include "runtime.cuh"
#include "curve.cuh"
#include <cstring>

namespace halo {

struct Bits256 { u32 w[8]; };

template <int C>
struct Host {
    static constexpr int S = CurveTraits<C>::SCALAR, B = CurveTraits<C>::BASE;
    typedef Fe<S> SF;
    typedef XYZZ<C> Pt;

    static Pt from_words(const uint64_t* w) { Jac<C> j; memcpy(&j, w, 96); return jac_to_xyzz<C>(j); }
    static void to_words(const Pt& p, uint64_t* w) { Jac<C> j = xyzz_to_jac<C>(p); memcpy(w, &j, 96); }
    static Pt from_affine_words(const uint64_t* w) { Affine<C> a; memcpy(&a, w, 64); return xyzz_from_affine<C>(a); }
    static SF scalar(const uint64_t* w) { SF s; memcpy(&s, w, 32); return s; }
    static void put_scalar(const SF& s, uint64_t* w) { memcpy(w, &s, 32); }
    static Bits256 bits(const SF& k) { SF c = fe_to_canonical<S>(k); Bits256 b; memcpy(b.w, c.l, 32); return b; }
    static Pt mul(const Pt& p, const SF& k) { Bits256 b = bits(k); return xyzz_mul_bits<C>(p, b.w, 255); }
    static Pt add(Pt a, const Pt& b) { xyzz_add<C>(a, b); return a; }
    static Pt neg(Pt a) { a.Y = fe_neg<B>(a.Y); return a; }
    static bool scalar_is_valid(const SF& s) {           // canonical limbs below the modulus
        for (int i = 7; i >= 0; --i) { u32 p = modulus_limb<S>(i); if (s.l[i] != p) return s.l[i] < p; }
        return false;
    }
    static bool scalar_words_valid(const uint64_t* w) { return scalar_is_valid(scalar(w)); }
    static bool base_is_valid(const Fe<B>& s) {
        for (int i = 7; i >= 0; --i) { u32 p = modulus_limb<B>(i); if (s.l[i] != p) return s.l[i] < p; }
        return false;
    }
    // What arkworks' validated deserialisation guarantees for a Projective<P> and the C ABI has to check itself:
    // every coordinate below the modulus and Y^2 = X^3 + 5 Z^6 (Z = 0 is the identity).  Both Pasta curves have prime
    // order, so on-curve implies in-subgroup.
    static bool point_words_valid(const uint64_t* w) {
        Jac<C> j; memcpy(&j, w, 96);
        if (!base_is_valid(j.X) || !base_is_valid(j.Y) || !base_is_valid(j.Z)) return false;
        if (fe_is_zero<B>(j.Z)) return true;
        const Fe<B> z2 = fe_sqr<B>(j.Z), z6 = fe_mul<B>(fe_sqr<B>(z2), z2);
        const Fe<B> rhs = fe_add<B>(fe_mul<B>(fe_sqr<B>(j.X), j.X), fe_mul<B>(fe_from_u64<B>(5), z6));
        return fe_eq<B>(fe_sqr<B>(j.Y), rhs);
    }
};

// msm.cu / ntt.cu / ipa.cu
template <int C> int msm_device(const void* d_bases, const void* d_scalars, size_t k, size_t nbatch, uint64_t* out_xyz,
                                ptrdiff_t base_stride);
template <int F> int ntt_device(Fe<F>* data, Fe<F>* out, int log_n, size_t batch, bool inverse, const Fe<F>* coset);
template <int C> int pcdl_open(const uint64_t* coeffs, size_t len, const uint64_t* C_xyz, size_t d, const uint64_t* z_w,
                               const uint64_t* v_w, const uint64_t* w_w, const uint64_t* hide_q, const uint64_t* hide_wbar,
                               uint64_t* out_proof, uint64_t* out_v, bool coeffs_resident, bool shard = false);
template <int C> int pcdl_check(const uint64_t* instance);
template <int C> int ensure_fixed_table(const void** tab, int* c, int* ncols, size_t* row_stride);   // commit-path table of pp.Gs (ipa.cu)
template <int C> int pcdl_open_prepare(size_t n);   // builds, on the main runtime, what an opening of size n would build lazily
template <int C> int acc_prover(const uint64_t* insts, size_t m, uint64_t* out_acc, bool shard_open = false);
template <int C> int acc_verifier(const uint64_t* insts, size_t m, const uint64_t* acc, bool sharded = false);
// dist.cu: the library's NCCL communicator (one process per GPU)
bool dist_on();
int dist_rank();
int dist_world();
int dist_allgather_dev(const void* d_send, void* d_recv, size_t bytes);
int dist_allgather_host(const void* h_send, void* h_recv_all, size_t bytes);
int dist_bcast_dev(void* d_buf, size_t bytes, int root);
int dist_bcast_host(void* h_buf, size_t bytes, int root);
bool dist_half_available();      // the communicator can be split into two halves (even world >= 4)
int dist_half_enter();           // from here on rank / world / collectives refer to the caller's half
void dist_half_leave();
bool dist_in_half();
template <int C> int dist_msm_device(const void* d_bases, const void* d_scalars, size_t k_local, size_t nbatch, uint64_t* out_xyz,
                                     ptrdiff_t base_stride = 0, bool from_host = false);

}  // namespace halo
