// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
//
// Variable-base MSM as reached by the reference through
// `Point::msm_unchecked(Gs, xs)` (crates/group/src/group.rs:48-56,
// crates/accumulation/src/pedersen.rs:21).  ark-ec 0.5.0's VariableBaseMSM is a
// signed-digit windowed Pippenger with window ~ ln(k)+2 parallelised with rayon
// over windows; the result is a group element, hence unique, so this
// restatement only has to be a correct MSM.  Conventions kept:
//   * pairs min(|bases|,|scalars|) items ("unchecked")
//   * scalars arrive in Montgomery form and are taken to canonical integers
//   * identity bases / zero scalars contribute nothing
// Threads split (point-chunk x window) tasks so that many-core hosts are used
// at least as well as rayon-over-windows would.
#pragma once
#include "curve.hpp"
#include <thread>
#include <atomic>
#include <cmath>

namespace orc {

static inline unsigned msm_window_bits(size_t k) {
    if (k < 32) return 3;
    return (unsigned)(std::log((double)k)) + 2;   // same rule of thumb as ark-ec
}

template <int CID>
Jac<CID> msm_serial_window(const Affine<CID>* bases, const int32_t* digits, size_t k, unsigned c) {
    size_t nb = (size_t)1 << (c - 1);
    std::vector<Jac<CID>> buckets(nb, Jac<CID>::identity());
    for (size_t i = 0; i < k; ++i) {
        int32_t d = digits[i];
        if (d == 0) continue;
        if (d > 0) buckets[d - 1] = buckets[d - 1].add_affine(bases[i]);
        else buckets[-d - 1] = buckets[-d - 1].add_affine(bases[i].neg());
    }
    Jac<CID> running = Jac<CID>::identity(), acc = Jac<CID>::identity();
    for (size_t b = nb; b-- > 0;) {
        running = running.add(buckets[b]);
        acc = acc.add(running);
    }
    return acc;
}

// scalars: k x 4 u64 Montgomery limbs of the curve's scalar field.
template <int CID>
Jac<CID> msm(const Affine<CID>* bases, const u64* scalars, size_t k, unsigned nthreads) {
    typedef typename Jac<CID>::S S;
    if (k == 0) return Jac<CID>::identity();
    if (nthreads == 0) nthreads = 1;
    unsigned c = msm_window_bits(k);
    unsigned nwin = (255 + c - 1) / c + 1;           // +1 absorbs the final signed-digit carry
    // signed digits, window-major
    std::vector<int32_t> digits((size_t)nwin * k);
    {
        auto work = [&](size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) {
                u64 w[4];
                S::from_raw(scalars + 4 * i).to_canonical(w);
                int carry = 0;
                for (unsigned j = 0; j < nwin; ++j) {
                    unsigned bit = j * c;
                    u64 v = 0;
                    if (bit < 256) {
                        unsigned limb = bit / 64, off = bit % 64;
                        v = w[limb] >> off;
                        if (off + c > 64 && limb + 1 < 4) v |= w[limb + 1] << (64 - off);
                        v &= ((u64)1 << c) - 1;
                    }
                    int64_t d = (int64_t)v + carry;
                    if (d > ((int64_t)1 << (c - 1))) { d -= (int64_t)1 << c; carry = 1; } else carry = 0;
                    digits[(size_t)j * k + i] = (int32_t)d;
                }
            }
        };
        std::vector<std::thread> th;
        size_t per = (k + nthreads - 1) / nthreads;
        for (unsigned t = 0; t < nthreads; ++t) {
            size_t lo = std::min(k, t * per), hi = std::min(k, lo + per);
            if (lo < hi) th.emplace_back(work, lo, hi);
        }
        for (auto& t : th) t.join();
    }
    unsigned nchunks = 1;
    while (nchunks * nwin < nthreads && k / (nchunks * 2) >= 4096) nchunks *= 2;
    size_t chunk = (k + nchunks - 1) / nchunks;
    std::vector<Jac<CID>> partial((size_t)nwin * nchunks, Jac<CID>::identity());
    std::atomic<size_t> next(0);
    auto worker = [&]() {
        for (;;) {
            size_t t = next.fetch_add(1);
            if (t >= (size_t)nwin * nchunks) break;
            unsigned j = (unsigned)(t / nchunks), ch = (unsigned)(t % nchunks);
            size_t lo = std::min(k, ch * chunk), hi = std::min(k, lo + chunk);
            partial[t] = msm_serial_window<CID>(bases + lo, digits.data() + (size_t)j * k + lo, hi - lo, c);
        }
    };
    {
        std::vector<std::thread> th;
        unsigned nt = std::min<size_t>(nthreads, (size_t)nwin * nchunks);
        for (unsigned t = 0; t < nt; ++t) th.emplace_back(worker);
        for (auto& t : th) t.join();
    }
    Jac<CID> total = Jac<CID>::identity();
    for (unsigned j = nwin; j-- > 0;) {
        for (unsigned b = 0; b < c; ++b) total = total.dbl();
        for (unsigned ch = 0; ch < nchunks; ++ch) total = total.add(partial[(size_t)j * nchunks + ch]);
    }
    return total;
}

// naive reference MSM (double-and-add per term) for cross-checking Pippenger on tiny inputs
template <int CID>
Jac<CID> msm_naive(const Affine<CID>* bases, const u64* scalars, size_t k) {
    typedef typename Jac<CID>::S S;
    Jac<CID> acc = Jac<CID>::identity();
    for (size_t i = 0; i < k; ++i)
        acc = acc.add(Jac<CID>::from_affine(bases[i]).mul(S::from_raw(scalars + 4 * i)));
    return acc;
}

}  // namespace orc
