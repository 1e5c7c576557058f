// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header).
// Moduli of the Pasta cycle as named by the reference
// (crates/group/src/lib.rs:8-9, SURVEY.md §0 "naming trap").
#include "field.hpp"

namespace orc {

static void mod_double(u64 x[4], const u64 p[4]) {
    u64 c = add4(x, x, x);
    if (c || geq4(x, p)) sub4(x, x, p);
}

static FieldParams make_params(const u64 p[4]) {
    FieldParams F;
    memcpy(F.p, p, 32);
    // -p^{-1} mod 2^64 by Newton iteration
    u64 x = 1;
    for (int i = 0; i < 7; ++i) x *= 2 - p[0] * x;
    F.inv = (u64)0 - x;
    u64 r[4] = {1, 0, 0, 0};
    for (int i = 0; i < 256; ++i) mod_double(r, p);
    memcpy(F.r, r, 32);
    for (int i = 0; i < 256; ++i) mod_double(r, p);
    memcpy(F.r2, r, 32);
    return F;
}

// halo Fp = ark_pallas::Fr ; halo Fq = ark_pallas::Fq
static const u64 P_FP[4] = {0x8c46eb2100000001ULL, 0x224698fc0994a8ddULL, 0x0ULL, 0x4000000000000000ULL};
static const u64 P_FQ[4] = {0x992d30ed00000001ULL, 0x224698fc094cf91bULL, 0x0ULL, 0x4000000000000000ULL};
const FieldParams FIELD_PARAMS[2] = {make_params(P_FP), make_params(P_FQ)};

}  // namespace orc
