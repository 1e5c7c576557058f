// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Radix-2 NTT / iNTT as reached by the reference through ark-poly 0.5.0
// `GeneralEvaluationDomain` (crates/group/src/poly.rs:11,56-64,133-139 and the
// direct calls at crates/plonk/src/plonk/protocol.rs:82-106,140-141,158-159,195).
// Conventions restated:
//   * omega_N = (5^((p-1)/2^32))^(2^(32-log N)); pinned by ivc/mod.rs:56,113
//   * forward: evals[i] = sum_j coeff[j] * omega^(i*j), natural order in and out
//   * inverse: includes the multiplication by N^{-1}
//   * coset (API addition, no reference call site): forward evaluates on
//     g*<omega>, inverse undoes it
#pragma once
#include "field.hpp"
#include <vector>
#include <thread>

namespace orc {

template <class F>
static void par_for(size_t n, unsigned nthreads, const F& fn) {
    if (nthreads <= 1 || n < 64) { fn(0, n); return; }
    std::vector<std::thread> th;
    size_t per = (n + nthreads - 1) / nthreads;
    for (unsigned t = 0; t < nthreads; ++t) {
        size_t lo = std::min(n, t * per), hi = std::min(n, lo + per);
        if (lo < hi) th.emplace_back([=, &fn]() { fn(lo, hi); });
    }
    for (auto& t : th) t.join();
}

template <int FID>
void ntt_inplace(Fe<FID>* a, unsigned log_n, bool inverse, const Fe<FID>* coset, unsigned nthreads) {
    typedef Fe<FID> F;
    size_t n = (size_t)1 << log_n;
    if (coset && !inverse) {
        // scale coefficient j by g^j
        par_for(n, nthreads, [&](size_t lo, size_t hi) {
            F g = coset->pow_u64(lo);
            for (size_t j = lo; j < hi; ++j) { a[j] = a[j] * g; g = g * *coset; }
        });
    }
    // bit reversal
    for (size_t i = 0; i < n; ++i) {
        size_t r = 0;
        for (unsigned b = 0; b < log_n; ++b) r |= ((i >> b) & 1) << (log_n - 1 - b);
        if (i < r) { F t = a[i]; a[i] = a[r]; a[r] = t; }
    }
    F w = root_of_unity<FID>(log_n);
    if (inverse) w = w.inv();
    // twiddle table w^k, k < n/2
    std::vector<F> tw(n / 2 ? n / 2 : 1);
    tw[0] = F::one();
    for (size_t k = 1; k < n / 2; ++k) tw[k] = tw[k - 1] * w;
    for (unsigned s = 1; s <= log_n; ++s) {
        size_t m = (size_t)1 << s, half = m >> 1, stride = n / m;
        par_for(n / 2, nthreads, [&](size_t lo, size_t hi) {
            for (size_t t = lo; t < hi; ++t) {
                size_t blk = t / half, j = t % half;
                size_t i0 = blk * m + j, i1 = i0 + half;
                F u = a[i0], v = a[i1] * tw[j * stride];
                a[i0] = u + v; a[i1] = u - v;
            }
        });
    }
    if (inverse) {
        F ninv = F::from_u64(n).inv();
        if (coset) {
            F gi = coset->inv();
            par_for(n, nthreads, [&](size_t lo, size_t hi) {
                F g = gi.pow_u64(lo) * ninv;
                for (size_t j = lo; j < hi; ++j) { a[j] = a[j] * g; g = g * gi; }
            });
        } else {
            par_for(n, nthreads, [&](size_t lo, size_t hi) {
                for (size_t j = lo; j < hi; ++j) a[j] = a[j] * ninv;
            });
        }
    }
}

// O(n^2) definition, for pinning ntt_inplace on tiny sizes
template <int FID>
void dft_naive(const Fe<FID>* in, Fe<FID>* out, unsigned log_n) {
    typedef Fe<FID> F;
    size_t n = (size_t)1 << log_n;
    F w = root_of_unity<FID>(log_n);
    F wi = F::one();
    for (size_t i = 0; i < n; ++i) {
        F acc = F::zero(), x = F::one();
        for (size_t j = 0; j < n; ++j) { acc = acc + in[j] * x; x = x * wi; }
        out[i] = acc;
        wi = wi * w;
    }
}

}  // namespace orc
