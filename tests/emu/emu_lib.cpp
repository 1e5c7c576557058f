// Host build of the DEVICE field/curve algorithms (halo_b200/csrc/{field,curve}.cuh) with the PTX carry-chain
// primitives emulated (-DHALO_EMULATE_PTX), or of the plain host path (-DEMU_PLAIN_HOST).  Lets the CPU test
// suite pin the exact limb-level algorithm the kernels run, without a GPU.  Test infrastructure only.
#ifndef EMU_PLAIN_HOST
#define HALO_EMULATE_PTX 1
#endif
#include "curve.cuh"
#include "transcript.cuh"
#include <cstring>
#include <cstddef>
using namespace halo;

template <int F> static void binop(int op, const u64* a, const u64* b, u64* o, size_t n) {
    for (size_t i = 0; i < n; ++i) {
        Fe<F> x, y, r; memcpy(x.l, a + 4 * i, 32); memcpy(y.l, b + 4 * i, 32);
        switch (op) {
            case 0: r = fe_mul<F>(x, y); break;
            case 1: r = fe_add<F>(x, y); break;
            case 2: r = fe_sub<F>(x, y); break;
            case 3: r = fe_inv<F>(x); break;
            case 4: r = fe_to_canonical<F>(x); break;
            case 5: r = fe_from_canonical<F>(x); break;
            case 7: r = fe_inv_safegcd<F>(x); break;
            default: r = fe_sqr<F>(x);
        }
        memcpy(o + 4 * i, r.l, 32);
    }
}
// acc = sum_i (+/-) P_i via xyzz_madd, starting from identity; then optionally add it to itself `dbls` times via
// xyzz_add (exercises the general add + doubling branch); output Jacobian
template <int C> static void madd_chain(const u64* pts, const int* neg, size_t n, int self_adds, u64* out_xyz) {
    XYZZ<C> acc = xyzz_identity<C>();
    for (size_t i = 0; i < n; ++i) { Affine<C> p; memcpy(&p, pts + 8 * i, 64); xyzz_madd<C>(acc, p, neg[i] != 0); }
    for (int k = 0; k < self_adds; ++k) { XYZZ<C> t = acc; xyzz_add<C>(acc, t); }
    Jac<C> j = xyzz_to_jac<C>(acc);
    memcpy(out_xyz, &j, 96);
}
template <int C> static void add2(const u64* a_xyz, const u64* b_xyz, u64* out_xyz, u64* out_xy) {
    Jac<C> a, b; memcpy(&a, a_xyz, 96); memcpy(&b, b_xyz, 96);
    XYZZ<C> x = jac_to_xyzz<C>(a), y = jac_to_xyzz<C>(b);
    xyzz_add<C>(x, y);
    Jac<C> j = xyzz_to_jac<C>(x); memcpy(out_xyz, &j, 96);
    Affine<C> af = xyzz_to_affine<C>(x); memcpy(out_xy, &af, 64);
}
template <int C> static void mulk(const u64* a_xyz, const u64* k_canon, u64* out_xyz) {
    Jac<C> a; memcpy(&a, a_xyz, 96);
    XYZZ<C> r = xyzz_mul_bits<C>(jac_to_xyzz<C>(a), (const u32*)k_canon);
    Jac<C> j = xyzz_to_jac<C>(r); memcpy(out_xyz, &j, 96);
}
template <int F> static void sponge_hash(const u64* in, size_t n, u64* out) {
    PoseidonHost<F> sp;
    for (size_t i = 0; i < n; ++i) { Fe<F> x; memcpy(x.l, in + 4 * i, 32); sp.absorb(x); }
    Fe<F> r = sp.squeeze();
    memcpy(out, r.l, 32);
}
template <int F> static void dot3(const u64* a, const u64* b, u64* o, size_t n) {
    for (size_t i = 0; i + 3 <= n; i += 3) {
        Fe<F> x[3], y[3];
        for (int k = 0; k < 3; ++k) { memcpy(x[k].l, a + 4 * (i + k), 32); memcpy(y[k].l, b + 4 * (i + k), 32); }
        Fe<F> r = fe_dot3<F>(x[0], y[0], x[1], y[1], x[2], y[2]);
        memcpy(o + 4 * (i / 3), r.l, 32);
    }
}
extern "C" {
// the product's host sponge (transcript.cuh): absorb n elements, squeeze one
void emu_poseidon_hash(int field, const u64* in, size_t n, u64* out) {
    if (field == 0) sponge_hash<0>(in, n, out); else sponge_hash<1>(in, n, out);
}
// out[i] = a[3i]*b[3i] + a[3i+1]*b[3i+1] + a[3i+2]*b[3i+2]
void emu_dot3(int field, const u64* a, const u64* b, u64* o, size_t n) {
    if (field == 0) dot3<0>(a, b, o, n); else dot3<1>(a, b, o, n);
}
void emu_field_op(int field, int op, const u64* a, const u64* b, u64* o, size_t n) {
    if (field == 0) binop<0>(op, a, b, o, n); else binop<1>(op, a, b, o, n);
}
void emu_madd_chain(int curve, const u64* pts, const int* neg, size_t n, int self_adds, u64* out_xyz) {
    if (curve == 0) madd_chain<0>(pts, neg, n, self_adds, out_xyz); else madd_chain<1>(pts, neg, n, self_adds, out_xyz);
}
void emu_add(int curve, const u64* a, const u64* b, u64* out_xyz, u64* out_xy) {
    if (curve == 0) add2<0>(a, b, out_xyz, out_xy); else add2<1>(a, b, out_xyz, out_xy);
}
void emu_mul(int curve, const u64* a, const u64* k, u64* out_xyz) {
    if (curve == 0) mulk<0>(a, k, out_xyz); else mulk<1>(a, k, out_xyz);
}
}
